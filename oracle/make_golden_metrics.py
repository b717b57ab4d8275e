#!/usr/bin/env python
"""Generate tests/golden/metric_reward_cases.npz by running the REFERENCE's own reward functions with
the BLEU / ROUGE-L scorers.  TEST INFRASTRUCTURE.

Runs only in the build container (needs ``/root/reference``).  ``utils.get_cst_reward`` /
``utils.get_self_critical_reward`` are imported and run unmodified (``/root/reference/utils.py:169-324``);
the scorer classes they are handed are the restated ones of ``oracle/bleu_rouge_oracle.py``, injected
under the names ``utils.py:10-17`` imports so that the ``isinstance(bcmr_scorer, Bleu)`` branches
(``utils.py:211-213, 277-279``: keep BLEU-4 only) and the list -> array conversion (``utils.py:216-217,
281-282``) are the reference's own.  The fixtures therefore pin the dict protocol, the BLEU-4 selection,
the candidate -> video map and the baselines against the reference, and the scorer arithmetic against
its restatement (absent dependency, see ``bleu_rouge_oracle.py``).

Usage:  python oracle/make_golden_metrics.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import bleu_rouge_oracle  # noqa: E402
from oracle.make_golden import EDGE_ROWS, import_reference  # noqa: E402
from cst_captioning_b200 import synth  # noqa: E402


def main():
    ref_utils, _ = import_reference()
    cases = {}
    scorers = {'bleu': bleu_rouge_oracle.Bleu(), 'rouge': bleu_rouge_oracle.Rouge()}

    # self-critical, 5 videos x 20 samples + greedy rows (config-3 shape), both use_eos settings
    b = synth.make_batch(5, 20, 20, seed=302, greedy=True)
    gts = [b['refs'][v].astype(np.int64) for v in range(5)]
    cases['sc/refs'], cases['sc/sampled'], cases['sc/greedy'] = b['refs'], b['sampled'], b['greedy']
    for name, scorer in scorers.items():
        for ue in (0, 1):
            rew, m, g = ref_utils.get_self_critical_reward(
                b['sampled'].astype(np.int64), b['greedy'].astype(np.int64), gts, scorer, expand_feat=1,
                seq_per_img=20, use_eos=ue)
            cases['sc/%s/eos%d/rewards' % (name, ue)] = rew[:, 0].copy()
            cases['sc/%s/eos%d/m_score' % (name, ue)] = np.float64(m)
            cases['sc/%s/eos%d/b_score' % (name, ue)] = np.float64(g)

    # CST with the GT-consensus baseline (config-4 shape) and the model-sample baseline (config 5)
    b = synth.make_batch(6, 20, 20, seed=303)
    gts = [b['refs'][v].astype(np.int64) for v in range(6)]
    bcmr = np.random.default_rng(6).random((6, 20))
    cases['cst/refs'], cases['cst/sampled'], cases['cst/bcmrscores'] = b['refs'], b['sampled'], bcmr
    for name, scorer in scorers.items():
        for tag, kw in (('gt_k7', dict(bcmrscores=bcmr, scb_captions=7, scb_baseline=1, use_mixer=1)),
                        ('star_k20', dict(scb_captions=20, scb_baseline=2)),
                        ('star_k3', dict(scb_captions=3, scb_baseline=2)),
                        ('none', dict(scb_captions=0))):
            rew, m, bs = ref_utils.get_cst_reward(b['sampled'].astype(np.int64), gts, scorer, expand_feat=1,
                                                  seq_per_img=20, use_eos=1, **kw)
            cases['cst/%s/%s/rewards' % (name, tag)] = rew[:, 0].copy()
            cases['cst/%s/%s/m_score' % (name, tag)] = np.float64(m)
            cases['cst/%s/%s/b_score' % (name, tag)] = np.float64(bs)

    # small vocabulary (repeated n-grams: clipping at work), ragged reference counts beyond 32,
    # one candidate per video (expand_feat = 0)
    rng = np.random.default_rng(304)
    many, many_nw = synth.make_refs(rng, 6, 45, vocab=40)
    keep = [45, 33, 1, 17, 40, 2]
    gts = [many[v, :keep[v]].astype(np.int64) for v in range(6)]
    samp = synth.corrupt(rng, many, many_nw, 1, 0.3, vocab=40)
    cases['ragged/refs'] = np.concatenate(gts).astype(np.int32)
    cases['ragged/ref_counts'] = np.array(keep, dtype=np.int64)
    cases['ragged/sampled'] = samp
    for name, scorer in scorers.items():
        rew, m, bs = ref_utils.get_cst_reward(samp.astype(np.int64), gts, scorer, expand_feat=0, seq_per_img=3,
                                              scb_captions=2, scb_baseline=2, use_eos=0)
        cases['ragged/%s/rewards' % name] = rew[:, 0].copy()
        cases['ragged/%s/m_score' % name] = np.float64(m)
        cases['ragged/%s/b_score' % name] = np.float64(bs)

    # hand-made edge rows (empty captions, interior <start>, no <end>, repeats) as candidates AND refs
    edge = np.array(EDGE_ROWS, dtype=np.int64)
    gts = [edge[:4], edge[4:7], edge[7:]]
    cand = np.concatenate([edge, edge[::-1]])[:18]     # 3 videos x 6
    cases['edge/refs'] = edge.astype(np.int32)
    cases['edge/ref_counts'] = np.array([4, 3, 3], dtype=np.int64)
    cases['edge/sampled'] = cand.astype(np.int32)
    for name, scorer in scorers.items():
        for ue in (0, 1):
            rew, m, _ = ref_utils.get_cst_reward(cand, gts, scorer, expand_feat=1, seq_per_img=6, scb_captions=0,
                                                 use_eos=ue)
            cases['edge/%s/eos%d/rewards' % (name, ue)] = rew[:, 0].copy()
            cases['edge/%s/eos%d/m_score' % (name, ue)] = np.float64(m)

    out = os.path.join(ROOT, 'tests', 'golden', 'metric_reward_cases.npz')
    np.savez_compressed(out, **cases)
    print('wrote', out, len(cases), 'arrays')


if __name__ == '__main__':
    main()
