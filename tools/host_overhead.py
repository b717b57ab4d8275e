import os, sys, time
sys.path.insert(0, '/root/repo' if os.path.isdir('/root/repo/tests') else os.getcwd())
import numpy as np, torch
import cst_captioning_b200 as cst
from cst_captioning_b200 import synth
dev = torch.device('cuda:0')
rng = np.random.default_rng(1)
corpus, _ = synth.make_refs(rng, 500, 20)
scorer = cst.CiderD(df=cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1))
b = synth.make_batch(64, 20, 20, seed=12, greedy=True)
c, gr, r = (torch.from_numpy(b[k]).to(dev) for k in ('sampled', 'greedy', 'refs'))
rs = cst.RefSet(r.reshape(-1, 30), per_video=20)
for name, fn in (('tensor refs', lambda: cst.get_self_critical_reward(c, gr, r, scorer, expand_feat=1, seq_per_img=20, use_eos=1)),
                 ('RefSet refs', lambda: cst.get_self_critical_reward(c, gr, rs, scorer, expand_feat=1, seq_per_img=20, use_eos=1))):
    for _ in range(50): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(2000): fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print('%s: host %.1f us per call (issue only), %.1f us per call incl. drain' % (name, (t1 - t0) / 2000 * 1e6, (t2 - t0) / 2000 * 1e6))

# where the host time of one call goes
import cProfile, pstats, io
fn = lambda: cst.get_self_critical_reward(c, gr, rs, scorer, expand_feat=1, seq_per_img=20, use_eos=1)
pr = cProfile.Profile()
pr.enable()
for _ in range(2000):
    fn()
pr.disable()
torch.cuda.synchronize()
out = io.StringIO()
pstats.Stats(pr, stream=out).sort_stats('tottime').print_stats(14)
print('\n'.join(l[:150] for l in out.getvalue().splitlines()[:30]))
