#!/usr/bin/env python
"""Developer tool: k_bleu / k_rouge alone on the headline batch (65,536 videos x 20 captions vs 20 refs),
CUDA events around `iters` launches each; one JSON line.  `--once`: a single launch of each after a
warm-up (what an ncu capture wants)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import __graft_entry__ as g
g.build()
import cst_captioning_b200 as cst
from cst_captioning_b200 import synth

once = '--once' in sys.argv
nums = [a for a in sys.argv[1:] if a.isdigit()]
n_videos = int(nums[0]) if nums else 65536
iters = 1 if once else 30
dev = torch.device('cuda:0')
b = synth.make_batch(n_videos, 20, 20, seed=1000)
cand = torch.from_numpy(b['sampled']).to(dev)
greedy = torch.from_numpy(synth.make_batch(n_videos, 20, 20, seed=1000, greedy=True)['greedy']).to(dev) if '--greedy' in sys.argv else None
refs = cst.RefSet(torch.from_numpy(b['refs']).to(dev).reshape(n_videos * 20, -1), per_video=20)
out = {'videos': n_videos, 'captions': n_videos * 20}
for name, sc in (('bleu4', cst.Bleu(device=dev)), ('rouge_l', cst.Rouge(device=dev))):
    for _ in range(2):
        sc.score_rows(cand, refs, 20, use_eos=1, greedy=greedy)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        s = sc.score_rows(cand, refs, 20, use_eos=1, greedy=greedy)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    out[name] = {'ms': ms, 'captions_per_s': n_videos * 20 * (2 if greedy is not None else 1) / ms * 1e3,
                 'mean_score': float(s.mean())}
print(json.dumps(out))
