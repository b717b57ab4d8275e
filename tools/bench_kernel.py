#!/usr/bin/env python
"""Developer tool: device-resident timing of the bench workload's reward call only.

    [CST_REWARD_LIB=...] python tools/bench_kernel.py [--videos N] [--steps K] [--int64] [--mode scb*|greedy|scb]
Prints one JSON line: ms per step, captions/s, m_score (a checksum to compare builds with).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cst_captioning_b200 as cst  # noqa: E402
from cst_captioning_b200 import synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--videos', type=int, default=65536)
ap.add_argument('--steps', type=int, default=20)
ap.add_argument('--int64', action='store_true')
ap.add_argument('--mode', default='scb*', choices=['scb*', 'greedy', 'scb'])
args = ap.parse_args()
dev = torch.device('cuda:0')
rng = np.random.default_rng(synth.SEED)
corpus, _ = synth.make_refs(rng, 6513, 20)
table = cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1)
scorer = cst.CiderD(df=table)
batch = synth.make_batch(args.videos, 20, 20, seed=1000, greedy=args.mode == 'greedy')
dt = torch.int64 if args.int64 else torch.int32
cand = torch.from_numpy(batch['sampled']).to(dev, dt)
refs = cst.RefSet(torch.from_numpy(batch['refs']).to(dev, dt).reshape(args.videos * 20, 30), per_video=20)
if args.mode == 'greedy':
    greedy = torch.from_numpy(batch['greedy']).to(dev, dt)

    def step():
        return cst.get_self_critical_reward(cand, greedy, refs, scorer, expand_feat=1, seq_per_img=20, use_eos=1)
elif args.mode == 'scb':
    bcmr = torch.rand(args.videos, 20, dtype=torch.float64, device=dev)

    def step():
        return cst.get_cst_reward(cand, refs, scorer, bcmrscores=bcmr, expand_feat=1, seq_per_img=20,
                                  scb_captions=20, scb_baseline=1, use_eos=1, use_mixer=1)
else:
    def step():
        return cst.get_cst_reward(cand, refs, scorer, expand_feat=1, seq_per_img=20, scb_captions=20,
                                  scb_baseline=2, use_eos=1)
for _ in range(3):
    out = step()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(args.steps):
    out = step()
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / args.steps
print(json.dumps({'lib': os.path.basename(cst.LIB_PATH),
                  'mode': args.mode, 'int64': args.int64, 'videos': args.videos, 'ms_per_step': ms,
                  'captions_per_s': args.videos * 20 / ms * 1e3, 'm_score': float(out[1]), 'b_score': float(out[2]),
                  'reward_sum_abs': float(out[0][:, 0].double().abs().sum())}))
