"""The five BASELINE.json configs on one B200 (device-resident inputs, CUDA events)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as g
g.build()
import cst_captioning_b200 as cst
from cst_captioning_b200 import synth
from cst_captioning_b200.consensus import consensus_scores

dev = torch.device('cuda:0')
rng = np.random.default_rng(synth.SEED)
corpus, _ = synth.make_refs(rng, 6513, 20)
d_corpus = torch.from_numpy(corpus).to(dev)
t0 = time.perf_counter()
table = cst.DfTable.from_rows(d_corpus, use_eos=1)
torch.cuda.synchronize()
df_build_s = time.perf_counter() - t0
scorer = cst.CiderD(df=table)


def timed(fn, reps):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))


out = {'df_table_build_s (6513x20 corpus, device, incl. allocation)': df_build_s, 'df_ngrams': len(table)}

b = synth.make_batch(64, 1, 20, seed=11)
c, r = torch.from_numpy(b['sampled']).to(dev), torch.from_numpy(b['refs']).to(dev)
ms = timed(lambda: cst.get_cst_reward(c, r, scorer, expand_feat=0, seq_per_img=1, scb_captions=0, use_eos=1), 200)
out['config1: 64x1 vs 20 refs, no baseline'] = {'ms_per_call': ms, 'captions_per_s': 64 / ms * 1e3}

t0 = time.perf_counter()
cons = consensus_scores(d_corpus)
torch.cuda.synchronize()
s = time.perf_counter() - t0
out['config2: GT consensus 6513x20 leave-one-out, corpus df per pass'] = {'seconds': s, 'captions_per_s': 6513 * 20 / s}

b = synth.make_batch(64, 20, 20, seed=12, greedy=True)
c, gr, r = (torch.from_numpy(b[k]).to(dev) for k in ('sampled', 'greedy', 'refs'))
ms = timed(lambda: cst.get_self_critical_reward(c, gr, r, scorer, expand_feat=1, seq_per_img=20, use_eos=1), 200)
out['config3: 1280 sampled + 1280 greedy, reward = sample - greedy'] = {'ms_per_call': ms, 'captions_per_s': 2560 / ms * 1e3}
# the same call captured once in a CUDA graph and replayed (what a launch-bound training loop would do)
side = torch.cuda.Stream(dev)
with torch.cuda.stream(side):
    cst.get_self_critical_reward(c, gr, r, scorer, expand_feat=1, seq_per_img=20, use_eos=1)
side.synchronize()
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph, stream=side):
    keep = cst.get_self_critical_reward(c, gr, r, scorer, expand_feat=1, seq_per_img=20, use_eos=1)
ms = timed(graph.replay, 200)
out['config3, CUDA-graph replay'] = {'ms_per_call': ms, 'captions_per_s': 2560 / ms * 1e3}

# config 3 with the references staged once (CookedRefs): the call scores from the images
cooked = cst.CookedRefs.cook(r, scorer, use_eos=1)
ms = timed(lambda: cst.get_self_critical_reward(c, gr, cooked, scorer, expand_feat=1, seq_per_img=20, use_eos=1), 200)
out['config3, staged references'] = {'ms_per_call': ms, 'captions_per_s': 2560 / ms * 1e3}

b = synth.make_batch(512, 20, 20, seed=13)
c, r = torch.from_numpy(b['sampled']).to(dev), torch.from_numpy(b['refs']).to(dev)
bc = torch.rand(512, 20, dtype=torch.float64, device=dev)
ms = timed(lambda: cst.get_cst_reward(c, r, scorer, bcmrscores=bc, expand_feat=1, seq_per_img=20, scb_captions=20,
                                      scb_baseline=1, use_eos=1, use_mixer=1), 200)
out['config4: SCB (GT baseline) batch 512'] = {'ms_per_call': ms, 'captions_per_s': 10240 / ms * 1e3}

b = synth.make_batch(65536, 20, 20, seed=1000)
c = torch.from_numpy(b['sampled']).to(dev)
r = cst.RefSet(torch.from_numpy(b['refs']).to(dev).reshape(-1, 30), per_video=20)
ms = timed(lambda: cst.get_cst_reward(c, r, scorer, expand_feat=1, seq_per_img=20, scb_captions=20, scb_baseline=2,
                                      use_eos=1), 30)
out['config5: SCB* 65536 videos x 20'] = {'ms_per_call': ms, 'captions_per_s': 1310720 / ms * 1e3}
print(json.dumps(out, indent=1))
