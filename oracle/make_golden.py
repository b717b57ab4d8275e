#!/usr/bin/env python
"""Generate tests/golden/* by running the REFERENCE's own code.  TEST INFRASTRUCTURE.

Runs only in the build container (needs ``/root/reference``; the GPU box has
none).  What is reference code here and what is not:

* ``compute_ciderdf.precook / build_dict / compute_doc_freq`` -- imported and
  run unmodified (``/root/reference/compute_ciderdf.py:19-93``).
* ``utils.array_to_str / get_cst_reward / get_self_critical_reward`` --
  imported and run unmodified (``/root/reference/utils.py:135-324``).  The
  import needs stand-ins for the absent third-party packages
  (``utils.py:8-17``); the only one that matters is ``CiderD``, for which the
  restated scorer of ``oracle/ciderd_oracle.py`` is injected so that the
  ``isinstance`` dispatch at ``utils.py:194,260`` fires.
* the scorer arithmetic itself -- restated, NOT reference code (absent
  dependency, see ``ciderd_oracle.py``).  Fixtures therefore pin the
  canonicalisation, n-gram extraction, df and baseline logic against the
  reference, and the scorer against its restatement.

Usage:  python oracle/make_golden.py            (writes tests/golden/)
"""
import json
import os
import pickle
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = '/root/reference'
sys.path.insert(0, ROOT)

from oracle import ciderd_oracle  # noqa: E402
from cst_captioning_b200 import synth  # noqa: E402


def import_reference():
    """-> (utils module, compute_ciderdf module) of the unmodified reference."""
    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class _Absent:  # Meteor/COCO are never instantiated on the path
        pass
    from oracle import bleu_rouge_oracle

    stub('pyciderevalcap')
    stub('pyciderevalcap.ciderD')
    stub('pyciderevalcap.ciderD.ciderD', CiderD=ciderd_oracle.CiderD)
    stub('pycocotools')
    stub('pycocotools.coco', COCO=_Absent)
    stub('pycocoevalcap')
    stub('pycocoevalcap.eval', COCOEvalCap=_Absent)
    # (the restated Bleu / Rouge, so that utils.py:211,277 `isinstance(bcmr_scorer, Bleu)` fires for
    # oracle/make_golden_metrics.py; the CIDEr fixtures below never touch them)
    for sub, cls, impl in (('bleu', 'Bleu', bleu_rouge_oracle.Bleu), ('rouge', 'Rouge', bleu_rouge_oracle.Rouge),
                           ('meteor', 'Meteor', _Absent)):
        stub('pycocoevalcap.' + sub)
        stub('pycocoevalcap.%s.%s' % (sub, sub), **{cls: impl})
    sys.path.insert(0, REF)
    import utils as ref_utils
    import compute_ciderdf as ref_df
    return ref_utils, ref_df


def reference_df(ref_df, refs, n_words, vocab):
    """df of a synthetic corpus through the reference's build_dict, fed the
    same structure ``compute_ciderdf.main`` builds (words + <end>, no <start>)."""
    names = ['<end>', '<start>', '<unk>'] + ['w%d' % i for i in range(3, vocab)]
    wtoi = {w: i for i, w in enumerate(names)}
    videos = []
    for v in range(refs.shape[0]):
        caps = []
        for r in range(refs.shape[1]):
            ids = refs[v, r, 1:1 + int(n_words[v, r])].tolist()
            caps.append([names[i] for i in ids] + ['<end>'])
        videos.append({'final_captions': caps})
    _, by_id, count = ref_df.build_dict(videos, wtoi)
    return by_id, count


def df_to_arrays(df):
    """dict{tuple[str]: float} -> (ids int32 (n,4) padded with -1, df float64)."""
    keys = sorted(df.keys(), key=lambda g: (len(g), tuple(int(t) for t in g)))
    ids = -np.ones((len(keys), 4), dtype=np.int32)
    val = np.zeros(len(keys), dtype=np.float64)
    for i, g in enumerate(keys):
        ids[i, :len(g)] = [int(t) for t in g]
        val[i] = df[g]
    return ids, val


EDGE_ROWS = [
    [1, 5, 7, 1, 9, 0, 4, 0],          # interior <start>, tokens after <end>
    [0, 4, 5, 6, 0, 0, 0, 0],          # empty caption
    [1, 1, 1, 1, 1, 1, 1, 1],          # only <start>
    [1, 0, 0, 0, 0, 0, 0, 0],          # <start> then <end>
    [5, 5, 5, 5, 5, 5, 5, 5],          # repeated token, no <end>
    [3, 4, 5, 6, 7, 8, 9, 10],         # full width, no <end>
    [7, 0, 0, 0, 0, 0, 0, 0],          # single word
    [9, 8, 1, 0, 3, 3, 3, 3],          # <start> right before <end>
    [2, 2, 0, 2, 2, 0, 0, 0],          # <unk>s
    [1, 12, 13, 12, 13, 12, 13, 0],    # repeated bigrams
]


def main():
    out_dir = os.path.join(ROOT, 'tests', 'golden')
    os.makedirs(out_dir, exist_ok=True)
    ref_utils, ref_df = import_reference()

    # ---- 1. canonicalisation + n-gram extraction truth tables -------------
    rng = np.random.default_rng(7)
    rows = [list(r) for r in EDGE_ROWS]
    for _ in range(40):
        w = int(rng.integers(1, 31))
        rows.append(rng.choice([0, 1, 2, 3, 4, 5, 6, 40, 41], size=w,
                               p=[.1, .08, .05, .2, .15, .12, .1, .1, .1]).tolist())
    canon = []
    for r in rows:
        arr = np.array(r, dtype=np.int64)
        entry = {'row': r}
        for ue in (0, 1):
            s = ref_utils.array_to_str(arr, ue)
            cooked = ref_df.precook(s)
            entry['str%d' % ue] = s
            entry['ngrams%d' % ue] = [[[int(t) for t in g], int(c)]
                                      for g, c in cooked.items()]
        canon.append(entry)
    with open(os.path.join(out_dir, 'canon_cases.json'), 'w') as fh:
        json.dump(canon, fh)

    # ---- 2. df through the reference's compute_ciderdf ---------------------
    cases = {}
    rng = np.random.default_rng(synth.SEED)
    df_refs, df_nw = synth.make_refs(rng, 60, 20)
    df_big, n_big = reference_df(ref_df, df_refs, df_nw, synth.VOCAB)
    ids, val = df_to_arrays(df_big)
    np.savez_compressed(os.path.join(out_dir, 'df_v10k.npz'), refs=df_refs,
                        n_words=df_nw, ngram_ids=ids, df=val, ref_len=n_big)

    # small vocabulary: many repeated n-grams, df > 1 at every order
    rng = np.random.default_rng(11)
    sm_refs, sm_nw = synth.make_refs(rng, 40, 6, vocab=40)
    df_small, n_small = reference_df(ref_df, sm_refs, sm_nw, 40)
    ids_s, val_s = df_to_arrays(df_small)
    np.savez_compressed(os.path.join(out_dir, 'df_v40.npz'), refs=sm_refs,
                        n_words=sm_nw, ngram_ids=ids_s, df=val_s, ref_len=n_small)

    # a genuine reference-format pickle (compute_ciderdf.py:123-129), tiny
    tiny_refs, tiny_nw = sm_refs[:8], sm_nw[:8]
    df_tiny, n_tiny = reference_df(ref_df, tiny_refs, tiny_nw, 40)
    with open(os.path.join(out_dir, 'df_tiny.pkl'), 'wb') as fh:
        pickle.dump({'document_frequency': df_tiny, 'ref_len': n_tiny}, fh,
                    protocol=pickle.HIGHEST_PROTOCOL)

    # ---- 3. reward functions of the reference ------------------------------
    scorer_big = ciderd_oracle.CiderD(df=(df_big, n_big))
    scorer_small = ciderd_oracle.CiderD(df=(df_small, n_small))

    def batch(n_videos, s, n_refs, seed, vocab, greedy=False):
        b = synth.make_batch(n_videos, s, n_refs, seed=seed, greedy=greedy,
                             vocab=vocab)
        gts = [b['refs'][v].astype(np.int64) for v in range(n_videos)]
        return b, gts

    def run_cst(tag, scorer, b, gts, **kw):
        res = b['sampled'].astype(np.int64)
        rew, m, bs = ref_utils.get_cst_reward(res, gts, scorer, **kw)
        cases[tag + '/rewards'] = rew[:, 0].copy()
        cases[tag + '/width'] = np.int64(rew.shape[1])
        cases[tag + '/m_score'] = np.float64(m)
        cases[tag + '/b_score'] = np.float64(bs)

    # config-1 shape: one sampled caption per video, no baseline
    b, gts = batch(16, 1, 20, 201, synth.VOCAB)
    cases['cfg1/refs'], cases['cfg1/sampled'] = b['refs'], b['sampled']
    run_cst('cfg1', scorer_big, b, gts, expand_feat=0, seq_per_img=1,
            scb_captions=0, use_eos=1)

    # config-3 shape: self-critical, sampled - greedy
    b, gts = batch(5, 20, 20, 202, synth.VOCAB, greedy=True)
    cases['cfg3/refs'], cases['cfg3/sampled'] = b['refs'], b['sampled']
    cases['cfg3/greedy'] = b['greedy']
    for ue in (0, 1):
        rew, m, g = ref_utils.get_self_critical_reward(
            b['sampled'].astype(np.int64), b['greedy'].astype(np.int64), gts,
            scorer_big, expand_feat=1, seq_per_img=20, use_eos=ue)
        cases['cfg3/eos%d/rewards' % ue] = rew[:, 0].copy()
        cases['cfg3/eos%d/m_score' % ue] = np.float64(m)
        cases['cfg3/eos%d/b_score' % ue] = np.float64(g)

    # config-4 shape: SCB with GT-consensus baseline
    b, gts = batch(6, 20, 20, 203, synth.VOCAB)
    bcmr = np.random.default_rng(5).random((6, 20)) * 3.0
    cases['cfg4/refs'], cases['cfg4/sampled'] = b['refs'], b['sampled']
    cases['cfg4/bcmrscores'] = bcmr
    for k in (20, 7, 1):
        run_cst('cfg4/k%d' % k, scorer_big, b, gts, bcmrscores=bcmr,
                expand_feat=1, seq_per_img=20, scb_captions=k, scb_baseline=1,
                use_eos=1, use_mixer=1)
    # WXE branch: weights are the precomputed scores themselves
    run_cst('cfg4/wxe', scorer_big, b, gts, bcmrscores=bcmr, expand_feat=1,
            seq_per_img=20, scb_captions=20, scb_baseline=1, use_eos=1,
            use_mixer=0)

    # config-5 shape: SCB* (model-sample baseline)
    b, gts = batch(6, 20, 20, 204, synth.VOCAB)
    cases['cfg5/refs'], cases['cfg5/sampled'] = b['refs'], b['sampled']
    for k in (20, 19, 3):
        for ue in (0, 1):
            run_cst('cfg5/k%d/eos%d' % (k, ue), scorer_big, b, gts,
                    expand_feat=1, seq_per_img=20, scb_captions=k,
                    scb_baseline=2, use_eos=ue)

    # small vocabulary, few refs, odd seq_per_img: heavy n-gram repetition
    b, gts = batch(9, 7, 6, 205, 40)
    cases['v40/refs'], cases['v40/sampled'] = b['refs'], b['sampled']
    for ue in (0, 1):
        run_cst('v40/eos%d' % ue, scorer_small, b, gts, expand_feat=1,
                seq_per_img=7, scb_captions=4, scb_baseline=2, use_eos=ue)

    # more than 32 refs per video and ragged ref counts (MSVD-like)
    rng = np.random.default_rng(206)
    many, many_nw = synth.make_refs(rng, 5, 45, vocab=40)
    keep = [45, 33, 1, 17, 40]
    gts = [many[v, :keep[v]].astype(np.int64) for v in range(5)]
    samp = synth.corrupt(rng, many, many_nw, 3, 0.3, vocab=40)
    cases['ragged/refs'] = np.concatenate([g for g in gts]).astype(np.int32)
    cases['ragged/ref_counts'] = np.array(keep, dtype=np.int64)
    cases['ragged/sampled'] = samp
    rew, m, bs = ref_utils.get_cst_reward(
        samp.astype(np.int64), gts, scorer_small, expand_feat=1, seq_per_img=3,
        scb_captions=2, scb_baseline=2, use_eos=1)
    cases['ragged/rewards'] = rew[:, 0].copy()
    cases['ragged/m_score'] = np.float64(m)
    cases['ragged/b_score'] = np.float64(bs)

    # hand-made edge rows as candidates AND as refs
    edge = np.array(EDGE_ROWS, dtype=np.int64)
    gts = [edge[:4], edge[4:7], edge[7:]]
    cand = np.concatenate([edge, edge[::-1]])[:18]     # 3 videos x 6
    cases['edge/refs'] = edge.astype(np.int32)
    cases['edge/ref_counts'] = np.array([4, 3, 3], dtype=np.int64)
    cases['edge/sampled'] = cand.astype(np.int32)
    for ue in (0, 1):
        rew, m, bs = ref_utils.get_cst_reward(
            cand, gts, scorer_small, expand_feat=1, seq_per_img=6,
            scb_captions=0, use_eos=ue)
        cases['edge/eos%d/rewards' % ue] = rew[:, 0].copy()
        cases['edge/eos%d/m_score' % ue] = np.float64(m)

    np.savez_compressed(os.path.join(out_dir, 'reward_cases.npz'), **cases)
    print('wrote', sorted(os.listdir(out_dir)))


if __name__ == '__main__':
    main()
