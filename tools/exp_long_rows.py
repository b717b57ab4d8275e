"""Developer experiment: how much do rows longer than 16 tokens (one-row-per-warp path) cost?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as g
g.build()
import cst_captioning_b200 as cst
from cst_captioning_b200 import synth
dev = torch.device('cuda:0')
rng = np.random.default_rng(synth.SEED)
corpus, _ = synth.make_refs(rng, 6513, 20)
scorer = cst.CiderD(df=cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1))
n = 32768
kw = dict(expand_feat=1, seq_per_img=20, scb_captions=20, scb_baseline=2, use_eos=1)
for cap in (28, 15, 12):
    b = synth.make_batch(n, 20, 20, seed=1000)
    refs, samp = b['refs'].copy(), b['sampled'].copy()
    refs[:, :, 1 + cap:] = 0          # cut every caption to at most `cap` words
    samp[:, cap:] = 0
    c = torch.from_numpy(samp).to(dev)
    r = cst.RefSet(torch.from_numpy(refs).to(dev).reshape(n * 20, 30), per_video=20)
    for _ in range(3):
        cst.get_cst_reward(c, r, scorer, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        cst.get_cst_reward(c, r, scorer, **kw)
    e1.record()
    torch.cuda.synchronize()
    words = (samp > 0).sum(1)
    print('max words %2d: %.3f ms per %d videos; mean words %.2f; rows > 15 words: %.2f%%'
          % (cap, e0.elapsed_time(e1) / 10, n, words.mean(), 100 * (words > 15).mean()))
