"""Small end-to-end run of every kernel for compute-sanitizer (memcheck / racecheck)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as g
g.build()
import cst_captioning_b200 as cst
from cst_captioning_b200 import synth
from cst_captioning_b200.consensus import consensus_scores
dev = torch.device('cuda:0')
rng = np.random.default_rng(1)
corpus, _ = synth.make_refs(rng, 64, 20)
scorer = cst.CiderD(df=cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1))
b = synth.make_batch(24, 20, 20, seed=5, greedy=True)
c = torch.from_numpy(b['sampled']).to(dev)
gr = torch.from_numpy(b['greedy']).to(dev)
r = torch.from_numpy(b['refs']).to(dev)
for kw in (dict(scb_captions=20, scb_baseline=2), dict(scb_captions=5, scb_baseline=1, bcmrscores=np.random.rand(24, 20)),
           dict(scb_captions=0)):
    cst.get_cst_reward(c, r, scorer, expand_feat=1, seq_per_img=20, use_eos=1, **kw)
cst.get_self_critical_reward(c, gr, r, scorer, expand_feat=1, seq_per_img=20, use_eos=1)
many = torch.from_numpy(synth.make_refs(rng, 3, 45, vocab=40)[0]).to(dev)
cst.get_cst_reward(c[:60], many, scorer, expand_feat=1, seq_per_img=20, scb_captions=3, scb_baseline=2, use_eos=1)
consensus_scores(r[:8, :6].contiguous())
seq = c.long()
lp = (-torch.rand(seq.shape, device=dev)).requires_grad_(True)
rew, _, _ = cst.get_cst_reward(c, r, scorer, expand_feat=1, seq_per_img=20, scb_captions=20, scb_baseline=2, use_eos=1)
cst.RewardCriterion()(seq, lp, rew).backward()
torch.cuda.synchronize()
print('sanitize run done')
