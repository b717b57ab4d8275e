"""Developer tool: average cycles a warp of k_score2 spends in each stage of a video.
Needs a library built with -DCST2_TIMING (tools/build_variant.sh T -DCST2_TIMING; CST_REWARD_LIB=...)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import cst_captioning_b200 as cst
from cst_captioning_b200 import synth, reward
dev = torch.device('cuda:0')
rng = np.random.default_rng(synth.SEED)
corpus, _ = synth.make_refs(rng, 6513, 20)
scorer = cst.CiderD(df=cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
b = synth.make_batch(n, 20, 20, seed=1000)
c = torch.from_numpy(b['sampled']).to(dev)
r = cst.RefSet(torch.from_numpy(b['refs']).to(dev).reshape(n * 20, 30), per_video=20)
kw = dict(expand_feat=1, seq_per_img=20, scb_captions=20, scb_baseline=2, use_eos=1)
for _ in range(3):
    cst.get_cst_reward(c, r, scorer, **kw)
torch.cuda.synchronize()
scratch = reward._scratch[(dev, torch.cuda.current_stream(dev).cuda_stream)]
t0 = scratch[64:64 + 128].clone().view(torch.int64).cpu().numpy()
cst.get_cst_reward(c, r, scorer, **kw)
torch.cuda.synchronize()
t1 = scratch[64:64 + 128].clone().view(torch.int64).cpu().numpy()
d = (t1 - t0).astype(np.float64)
warps = 8 * n
names = ['R load+canon', 'R barrier 1', 'R scan+emit', 'R barrier 2', 'R process', 'R barrier 3',
         'C load+canon', 'C barrier 1', 'C scan+emit', 'C barrier 2', 'C process', 'C barrier 3',
         'segment sums (both)', 'tail: scores, baseline', '-', '-']
for nm, v in zip(names, d / warps):
    print('%-22s %10.0f cycles per warp per video' % (nm, v))
print('total %10.0f' % (d.sum() / warps))
