"""The CPU oracle against the fixtures produced by the reference's own code
(oracle/make_golden.py).  No GPU."""
import json
import os

import numpy as np
import pytest

from oracle import ciderd_oracle as co
from oracle import reward_oracle as ro
from cst_captioning_b200 import synth

from conftest import GOLDEN


def test_canonicalisation_and_ngrams_match_reference():
    cases = json.load(open(os.path.join(GOLDEN, 'canon_cases.json')))
    assert len(cases) >= 50
    for c in cases:
        for ue in (0, 1):
            s = ro.row_to_str(c['row'], ue)
            assert s == c['str%d' % ue]
            got = [[[int(t) for t in g], int(n)] for g, n in co.ngram_counts(s).items()]
            assert got == c['ngrams%d' % ue]      # same keys, counts AND order


def test_survey_truth_table():
    # SURVEY.md §8c item (2)
    assert ro.row_to_str([1, 5, 7, 1, 9, 0, 4, 0], 0) == '5 7 9'
    assert ro.row_to_str([1, 5, 7, 1, 9, 0, 4, 0], 1) == '5 7 9 0'
    assert ro.row_to_str([0, 4], 0) == ''
    assert ro.row_to_str([0, 4], 1) == '0'


@pytest.mark.parametrize('name', ['df_v10k', 'df_v40'])
def test_document_frequency_matches_reference(name):
    z = np.load(os.path.join(GOLDEN, name + '.npz'))
    strings = synth.df_corpus_strings(z['refs'], z['n_words'])
    df, n = co.build_df_from_strings(strings)
    assert n == int(z['ref_len'])
    want = {tuple(str(t) for t in row if t >= 0): v
            for row, v in zip(z['ngram_ids'].tolist(), z['df'].tolist())}
    assert dict(df) == want


def test_reference_pickle_loads():
    table, n = co.load_df_pickle(os.path.join(GOLDEN, 'df_tiny.pkl'))
    assert n == 8 and len(table) > 100
    assert all(isinstance(k, tuple) and isinstance(k[0], str) for k in table)


def _gts(refs):
    return [refs[v].astype(np.int64) for v in range(refs.shape[0])]


def _split(rows, counts):
    out, at = [], 0
    for c in counts:
        out.append(rows[at:at + c].astype(np.int64))
        at += c
    return out


def test_cst_reward_matches_reference(golden_cases, golden_df_v10k, golden_df_v40):
    g = golden_cases
    big = co.CiderD(df=(golden_df_v10k[1], golden_df_v10k[2]))
    small = co.CiderD(df=(golden_df_v40[1], golden_df_v40[2]))

    def check(tag, rew, m, b):
        np.testing.assert_array_equal(rew[:, 0], g[tag + '/rewards'])
        assert m == g[tag + '/m_score']
        if tag + '/b_score' in g.files:
            assert b == g[tag + '/b_score']

    r = ro.get_cst_reward(g['cfg1/sampled'].astype(np.int64), _gts(g['cfg1/refs']), big,
                          expand_feat=0, seq_per_img=1, scb_captions=0, use_eos=1)
    check('cfg1', *r)
    assert r[0].shape == (16, 28)

    for k in (20, 7, 1):
        r = ro.get_cst_reward(g['cfg4/sampled'].astype(np.int64), _gts(g['cfg4/refs']), big,
                              bcmrscores=g['cfg4/bcmrscores'], expand_feat=1, seq_per_img=20,
                              scb_captions=k, scb_baseline=1, use_eos=1, use_mixer=1)
        check('cfg4/k%d' % k, *r)
    r = ro.get_cst_reward(g['cfg4/sampled'].astype(np.int64), _gts(g['cfg4/refs']), big,
                          bcmrscores=g['cfg4/bcmrscores'], expand_feat=1, seq_per_img=20,
                          scb_captions=20, scb_baseline=1, use_eos=1, use_mixer=0)
    check('cfg4/wxe', *r)

    for k in (20, 19, 3):
        for ue in (0, 1):
            r = ro.get_cst_reward(g['cfg5/sampled'].astype(np.int64), _gts(g['cfg5/refs']), big,
                                  expand_feat=1, seq_per_img=20, scb_captions=k,
                                  scb_baseline=2, use_eos=ue)
            check('cfg5/k%d/eos%d' % (k, ue), *r)

    for ue in (0, 1):
        r = ro.get_cst_reward(g['v40/sampled'].astype(np.int64), _gts(g['v40/refs']), small,
                              expand_feat=1, seq_per_img=7, scb_captions=4, scb_baseline=2,
                              use_eos=ue)
        check('v40/eos%d' % ue, *r)

    r = ro.get_cst_reward(g['ragged/sampled'].astype(np.int64),
                          _split(g['ragged/refs'], g['ragged/ref_counts']), small,
                          expand_feat=1, seq_per_img=3, scb_captions=2, scb_baseline=2, use_eos=1)
    check('ragged', *r)

    for ue in (0, 1):
        r = ro.get_cst_reward(g['edge/sampled'].astype(np.int64),
                              _split(g['edge/refs'], g['edge/ref_counts']), small,
                              expand_feat=1, seq_per_img=6, scb_captions=0, use_eos=ue)
        check('edge/eos%d' % ue, *r)

    with pytest.raises(ValueError):
        ro.get_cst_reward(g['cfg5/sampled'].astype(np.int64), _gts(g['cfg5/refs']), big,
                          expand_feat=1, seq_per_img=20, scb_captions=3, scb_baseline=3)


def test_self_critical_reward_matches_reference(golden_cases, golden_df_v10k):
    g = golden_cases
    big = co.CiderD(df=(golden_df_v10k[1], golden_df_v10k[2]))
    for ue in (0, 1):
        rew, m, b = ro.get_self_critical_reward(
            g['cfg3/sampled'].astype(np.int64), g['cfg3/greedy'].astype(np.int64),
            _gts(g['cfg3/refs']), big, expand_feat=1, seq_per_img=20, use_eos=ue)
        np.testing.assert_array_equal(rew[:, 0], g['cfg3/eos%d/rewards' % ue])
        assert m == g['cfg3/eos%d/m_score' % ue]
        assert b == g['cfg3/eos%d/b_score' % ue]


def test_scorer_hand_cases():
    """Values that can be checked by hand."""
    # two documents; 'a' in both (idf 0), the rest in one (idf log 2)
    df, n = co.build_df_from_strings([['3 4 5'], ['3 6']])
    sc = co.CiderD(df=(df, n))
    # hypothesis identical to its single ref: every order with non-zero norm is 1
    _, s = sc.compute_score({0: ['3 4 5']}, [{'image_id': 0, 'caption': ['3 4 5']}])
    # orders 1,2,3 have cosine 1; order 4 has no n-grams -> 0;  mean = 3/4, x10
    assert s[0] == pytest.approx(7.5, rel=1e-12)
    # disjoint hypothesis scores 0; empty hypothesis scores 0
    _, s = sc.compute_score({0: ['3 4 5']}, [{'image_id': 0, 'caption': ['7 8 9']}])
    assert s[0] == 0.0
    _, s = sc.compute_score({0: ['3 4 5']}, [{'image_id': 0, 'caption': ['']}])
    assert s[0] == 0.0
    # length penalty: same bag at order 1, bigram-count delta = 2
    _, s = sc.compute_score({0: ['4']}, [{'image_id': 0, 'caption': ['4 4 4']}])
    # unigram: vec_h = 3w, vec_r = w -> min*vr/(3w*w) = 1/3; penalty e^{-4/72}
    assert s[0] == pytest.approx(10 * (1 / 3) * np.exp(-4 / 72) / 4, rel=1e-12)


def test_corpus_mode_and_consensus_shape():
    rng = np.random.default_rng(3)
    refs, nw = synth.make_refs(rng, 12, 5, vocab=30)
    strings = [[ro.row_to_str(refs[v, r], 0) for r in range(5)] for v in range(12)]
    out = ro.consensus_scores(strings, seq_per_img=5)
    assert out.shape == (12, 5) and np.all(out >= 0) and out.max() > 0
    # permuting the refs of a video does not change a hypothesis' score
    sc = co.Cider()
    gts = {v: strings[v][1:] for v in range(12)}
    res = {v: [strings[v][0]] for v in range(12)}
    _, a = sc.compute_score(gts, res)
    gts2 = {v: strings[v][1:][::-1] for v in range(12)}
    _, b = sc.compute_score(gts2, res)
    np.testing.assert_allclose(a, b, rtol=1e-12)
