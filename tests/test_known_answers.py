"""Hand-derived CIDEr-D known answers, run through BOTH the CPU oracle and the CUDA path.

The scorer the reference imports (``/root/reference/utils.py:8-17``) is absent, so the oracle
in ``oracle/`` is a restatement.  These cases do not lean on it: every expected value below is
a closed form derived in the comments from the published CIDEr-D formula and evaluated with
``math`` only --

    score(h) = 10 / R * sum_r  1/4 * sum_{n=1..4}  [ sum_g min(h_g, r_g) * r_g / (|h|_n |r|_n) ] * exp(-d^2 / (2 * 6^2))

with  h_g = tf_h(g) * idf(g),  idf(g) = log(N) - log(max(1, df(g))),  |.|_n the Euclidean norm
over the n-grams of order n (an order with a zero norm on either side contributes 0),  and
d = (#bigrams of h) - (#bigrams of r)  the scorer's "length" difference.

Corpus of the cases: N = 8 documents, so with a = log 2:
    df 1 or unseen -> idf 3a,   df 2 -> 2a,   df 4 -> a,   df 8 -> 0.

The oracle half runs everywhere; the CUDA half (``-m gpu``) calls ``cst_cider_scores`` through the
C ABI on token rows and the string protocol ``CiderD.compute_score``.
"""
import math

import numpy as np
import pytest

from oracle import ciderd_oracle as co

A = math.log(2.0)
N_DOCS = 8
DF = {('3',): 2.0, ('4',): 4.0, ('6',): 8.0, ('0',): 8.0, ('6', '7'): 8.0, ('9',): 1.0}
#      idf 2a        idf a        idf 0        idf 0 (<end>)  idf 0           idf 3a (like unseen)


def g(d):
    """Gaussian length penalty for a bigram-count difference d (sigma = 6)."""
    return math.exp(-(d * d) / 72.0)


# name, use_eos, hypothesis row, reference rows, expected score (derivation in the comment)
CASES = [
    # every n-gram unseen (idf 3a): h = r on orders 1..3 -> cosine 1 each; order 4 has no n-gram -> 0
    ('identical_to_only_ref', 0, [10, 11, 12], [[10, 11, 12]], 10 * 3 / 4),
    # scores average over the refs: the disjoint second ref contributes 0
    ('multi_ref_average', 0, [10, 11, 12], [[10, 11, 12], [13, 14, 15]], 10 * (3 / 4) / 2),
    # clipping, tf_h = 3 > tf_r = 1 (idf w = 3a): order 1: min(3w, w) * w / (3w * w) = 1/3; the ref has no
    # bigram (norm 0) -> orders 2..4 give 0; d = 2 - 0
    ('clip_tf_h_gt_tf_r', 0, [10, 10, 10], [[10]], 10 / 4 * (1 / 3) * g(2)),
    # tf_h = 1 < tf_r = 3: min(w, 3w) * 3w / (w * 3w) = 1; d = 0 - 2
    ('clip_tf_h_lt_tf_r', 0, [10], [[10, 10, 10]], 10 / 4 * 1.0 * g(-2)),
    # d = -1: order 1: 2w^2 / (sqrt2 w * sqrt3 w) = 2/sqrt6; order 2: w^2 / (w * sqrt2 w) = 1/sqrt2
    ('gauss_minus_1', 0, [10, 11], [[10, 11, 12]], 10 / 4 * (2 / math.sqrt(6) + 1 / math.sqrt(2)) * g(-1)),
    # d = +1: the mirror image, same cosines
    ('gauss_plus_1', 0, [10, 11, 12], [[10, 11]], 10 / 4 * (2 / math.sqrt(6) + 1 / math.sqrt(2)) * g(1)),
    # d = -5: order 1: 2w^2 / (sqrt2 w * sqrt7 w) = 2/sqrt14; order 2: 1/sqrt6 (1 of the ref's 6 bigrams)
    ('gauss_minus_5', 0, [10, 11], [[10, 11, 12, 13, 14, 15, 16]],
     10 / 4 * (2 / math.sqrt(14) + 1 / math.sqrt(6)) * g(-5)),
    # d = +5
    ('gauss_plus_5', 0, [10, 11, 12, 13, 14, 15, 16], [[10, 11]],
     10 / 4 * (2 / math.sqrt(14) + 1 / math.sqrt(6)) * g(5)),
    # idf from df with the max(1, df) clamp: '3' -> 2a, '4' -> a, '5' unseen -> 3a.
    # order 1: h = {3: 2a, 5: 3a}, r = {3: 2a, 4: a}: (2a)^2 / (a sqrt13 * a sqrt5) = 4/sqrt65;
    # order 2: '3 5' vs '3 4': no common bigram -> 0
    ('idf_values_and_unseen_clamp', 0, [3, 5], [[3, 4]], 10 / 4 * 4 / math.sqrt(65)),
    # df 1 and "unseen" give the same idf 3a: '9' (df 1) against '5' (absent): cos of identical rows = 1,
    # and a row mixing both has norm a * 3 sqrt2: '9 5' vs '9': order 1: 9a^2 / (3 sqrt2 a * 3a) = 1/sqrt2; d = 1
    ('df1_equals_unseen', 0, [9, 5], [[9]], 10 / 4 * (1 / math.sqrt(2)) * g(1)),
    # an n-gram every document holds has idf 0: '6' alone -> both norms 0 -> nothing is divided, score 0
    ('idf_zero_only', 0, [6], [[6]], 0.0),
    # '6 7': unigram 6 -> 0, 7 -> 3a; bigram '6 7' has df 8 -> idf 0.  order 1: 9a^2 / (3a * 3a) = 1; order 2: 0
    ('idf_zero_mixed', 0, [6, 7], [[6, 7]], 10 / 4 * 1.0),
    # zero norm on the hypothesis side only: h = '6' (idf 0), r = '6 10': order 1 norm_h = 0 -> 0, no NaN
    ('zero_norm_one_side', 0, [6], [[6, 10]], 0.0),
    ('empty_hypothesis', 0, [], [[10, 11, 12]], 0.0),
    # use_eos = 0 drops the <end>: '10 11' vs '10 11' -> orders 1, 2 -> 10 * 2/4
    ('eos_dropped', 0, [10, 11, 0], [[10, 11, 0]], 10 * 2 / 4),
    # use_eos = 1 keeps it as the unigram '0' (df 8 -> idf 0) and inside the bigrams/trigram (unseen, 3a):
    # '10 11 0' vs itself: orders 1, 2, 3 have non-zero norms and cosine 1
    ('eos_kept', 1, [10, 11, 0], [[10, 11, 0]], 10 * 3 / 4),
    # use_eos = 1, '10 0' vs '10 11 0': order 1: h = {10: w, 0: 0}, r = {10: w, 11: w, 0: 0}: w^2 / (w * sqrt2 w)
    # = 1/sqrt2; order 2: '10 0' vs {'10 11', '11 0'}: nothing shared -> 0; d = 1 - 2
    ('eos_as_unigram', 1, [10, 0], [[10, 11, 0]], 10 / 4 * (1 / math.sqrt(2)) * g(-1)),
    # clipping on two orders: '10 10' vs '10 10 10 10': order 1: min(2w, 4w) * 4w / (2w * 4w) = 1;
    # order 2: min(w, 3w) * 3w / (w * 3w) = 1; order 3: the hypothesis has none; d = 1 - 3
    ('clip_two_orders', 0, [10, 10], [[10, 10, 10, 10]], 10 / 4 * 2.0 * g(-2)),
    # <start> (id 1) is dropped wherever it stands (utils.py:142-143): '1 10 1 11 12' scores like '10 11 12'
    ('start_token_skipped', 0, [1, 10, 1, 11, 12], [[1, 10, 11, 12]], 10 * 3 / 4),
    # two refs of different length: 0.75 from the identical one, gauss_plus_1's value from the short one
    ('two_refs_two_penalties', 0, [10, 11, 12], [[10, 11, 12], [10, 11]],
     10 / 2 * (3 / 4 + (2 / math.sqrt(6) + 1 / math.sqrt(2)) / 4 * g(1))),
]
assert len(CASES) >= 12


def _canon(ids, use_eos):
    """utils.array_to_str (utils.py:135-152) on a short list that may end in <end>."""
    out = []
    for t in ids:
        if t == 0:
            if use_eos:
                out.append('0')
            break
        if t == 1:
            continue
        out.append(str(t))
    return ' '.join(out)


@pytest.mark.parametrize('name,use_eos,hyp,refs,want', CASES, ids=[c[0] for c in CASES])
def test_oracle_reproduces_the_closed_form(name, use_eos, hyp, refs, want):
    scorer = co.CiderD(df=(DF, N_DOCS))
    gts = {0: [_canon(r, use_eos) for r in refs]}
    res = [{'image_id': 0, 'caption': [_canon(hyp, use_eos)]}]
    _, got = scorer.compute_score(gts, res)
    assert got[0] == pytest.approx(want, rel=1e-12, abs=1e-15)


def _rows(lists, width):
    out = np.zeros((len(lists), width), dtype=np.int32)       # 0 = <end> pads the rows like the reference's
    for i, ids in enumerate(lists):
        out[i, :len(ids)] = ids
    return out


@pytest.mark.gpu
def test_cuda_path_reproduces_the_closed_forms():
    """All cases in ONE launch of cst_cider_scores per use_eos (videos with ragged ref counts), and
    each case once more through the string protocol."""
    import torch
    import cst_captioning_b200 as cst
    dev = torch.device('cuda:0')
    scorer = cst.CiderD(df=(DF, N_DOCS), device=dev)
    for ue in (0, 1):
        cases = [c for c in CASES if c[1] == ue]
        cand = torch.from_numpy(_rows([c[2] for c in cases], 8)).to(dev)
        ref_rows = _rows([r for c in cases for r in c[3]], 8)
        offsets = np.concatenate([[0], np.cumsum([len(c[3]) for c in cases])])
        refs = cst.RefSet(torch.from_numpy(ref_rows).to(dev), offsets=offsets)
        got = scorer.score_rows(cand, refs, 1, use_eos=ue).cpu().numpy()
        for c, s in zip(cases, got):
            assert s == pytest.approx(c[4], rel=1e-12, abs=1e-15), c[0]
        # int64 rows (what the reference's model emits) give the same bits
        got64 = scorer.score_rows(cand.long(), cst.RefSet(refs.rows.long(), offsets=offsets), 1, use_eos=ue)
        np.testing.assert_array_equal(got64.cpu().numpy(), got)
        for c in cases:
            gts = {7: [_canon(r, ue) for r in c[3]]}
            _, s = scorer.compute_score(gts, [{'image_id': 7, 'caption': [_canon(c[2], ue)]}])
            assert s[0] == pytest.approx(c[4], rel=1e-12, abs=1e-15), c[0]


@pytest.mark.gpu
def test_cuda_reward_of_known_scores():
    """The closed-form scores through cst_reward: SCB* with k = 2 of 3 captions per video
    (utils.py:301-304, 311-312): baseline = mean of the two LOWEST scores of the row, own score included."""
    import torch
    import cst_captioning_b200 as cst
    dev = torch.device('cuda:0')
    scorer = cst.CiderD(df=(DF, N_DOCS), device=dev)
    hyps = [[10, 11, 12], [10, 11], [13]]                     # vs the single ref '10 11 12'
    s = np.array([7.5, 10 / 4 * (2 / math.sqrt(6) + 1 / math.sqrt(2)) * g(-1), 0.0])
    base = np.sort(s)[:2].mean()
    rew, m, b = cst.get_cst_reward(torch.from_numpy(_rows(hyps, 8)).to(dev),
                                   torch.from_numpy(_rows([[10, 11, 12]], 8)).to(dev).reshape(1, 1, 8), scorer,
                                   expand_feat=1, seq_per_img=3, scb_captions=2, scb_baseline=2, use_eos=0)
    np.testing.assert_allclose(rew[:, 0].double().cpu().numpy(), (s - base).astype(np.float32), rtol=1e-6)
    assert float(m) == pytest.approx(s.mean(), rel=1e-12)
    assert float(b) == pytest.approx(base, rel=1e-12)
