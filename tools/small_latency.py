"""Developer tool: per-call time of the fused reward kernel at small batch sizes
(CUDA events, median of 200 calls, device-resident inputs)."""
import os, sys, json
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


def timed(fn, reps=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3


out = {}
for n in (1, 8, 64, 148, 296, 592, 1184, 4096):
    for s in (1, 20):
        b = synth.make_batch(n, s, 20, seed=5)
        c = torch.from_numpy(b['sampled']).to(dev)
        r = cst.RefSet(torch.from_numpy(b['refs']).to(dev).reshape(n * 20, 30), per_video=20)
        kw = dict(expand_feat=1, seq_per_img=s, scb_captions=0, use_eos=1)
        us = timed(lambda: cst.get_cst_reward(c, r, scorer, **kw))
        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(dev)
        with torch.cuda.stream(side):
            cst.get_cst_reward(c, r, scorer, **kw)
        side.synchronize()
        with torch.cuda.graph(graph, stream=side):
            keep = cst.get_cst_reward(c, r, scorer, **kw)
        out['videos=%d cands=%d' % (n, s)] = {'us_per_call': round(us, 1), 'us_graph_replay': round(timed(graph.replay), 1)}
print(json.dumps(out, indent=1))
