"""BLEU-4 and ROUGE-L as reward scorers (``train.py:119-124``).

Three layers, as for CIDEr-D:

* hand-derived known answers (closed forms from the published formulas, evaluated with ``math`` only)
  through the CPU oracle and -- ``-m gpu`` -- the CUDA kernels, rows and strings;
* the fixtures of ``tests/golden/metric_reward_cases.npz``, which the REFERENCE's own
  ``utils.get_cst_reward`` / ``get_self_critical_reward`` produced with these scorers
  (``oracle/make_golden_metrics.py``): the oracle's reward functions and the CUDA path reproduce them;
* the CUDA kernels against the oracle on seeded random batches of every shape the kernels branch on.

BLEU per sentence (``bleu_scorer.py``, option 'closest'), with small = 1e-9, tiny = 1e-15:
    bleu_k = ( prod_{j<=k} (correct_j + tiny) / (guess_j + small) ) ** (1/k)  *  bp
    bp = exp(1 - 1/ratio) if ratio < 1 else 1,   ratio = (testlen + tiny) / (reflen + small)
    guess_j = max(testlen - j + 1, 0),  correct_j = sum over the candidate's j-grams of
    min(count in the candidate, largest count in any ONE reference),  reflen = the reference length
    closest to testlen (the shorter on a tie).
ROUGE-L (``rouge.py``): P = max_r LCS/len(cand), R = max_r LCS/len(ref_r), beta = 1.2,
    F = (1 + beta^2) P R / (R + beta^2 P), 0 if either maximum is 0; sentences cut by split(" ").
"""
import math
import os

import numpy as np
import pytest

from oracle import bleu_rouge_oracle as bo
from oracle import reward_oracle as ro

from conftest import GOLDEN

SMALL, TINY = 1e-9, 1e-15


def bleu_closed(correct, testlen, reflen):
    """BLEU-1..4 from hand-counted statistics."""
    ratio = (testlen + TINY) / (reflen + SMALL)
    bp = math.exp(1 - 1 / ratio) if ratio < 1 else 1.0
    out, prod = [], 1.0
    for k in range(4):
        prod *= (correct[k] + TINY) / (max(testlen - k, 0) + SMALL)
        out.append(prod ** (1.0 / (k + 1)) * bp)
    return out


# name, use_eos, candidate row, reference rows, (correct[4], testlen, reflen) counted by hand
BLEU_CASES = [
    # identical five words: every n-gram matches once; reflen 5 -> ratio just below 1 (small vs tiny)
    ('identical', 0, [10, 11, 12, 13, 14], [[10, 11, 12, 13, 14]], ([5, 4, 3, 2], 5, 5)),
    # a prefix of the reference: all its n-grams match, brevity penalty exp(1 - 4/2)
    ('brevity_penalty', 0, [10, 11], [[10, 11, 12, 13]], ([2, 1, 0, 0], 2, 4)),
    # clipping by the largest count in any ONE reference: '10' x3 vs refs '10' and '10 10':
    # unigram min(3, 2) = 2, bigram '10 10' min(2, 1) = 1, trigram min(1, 0) = 0; closest reflen to 3 is 2
    ('clip_to_max_over_refs', 0, [10, 10, 10], [[10], [10, 10]], ([2, 1, 0, 0], 3, 2)),
    # counts do not ADD over references: '10 10' vs refs '10', '10': unigram min(2, 1) = 1
    ('max_not_sum', 0, [10, 10], [[10], [10]], ([1, 0, 0, 0], 2, 1)),
    # reference lengths 2 and 4 are equally close to 3: the shorter wins -> ratio 1.5, no penalty
    ('closest_length_tie', 0, [10, 11, 12], [[10, 11], [10, 11, 12, 13]], ([3, 2, 1, 0], 3, 2)),
    # nothing in common: correct = 0 everywhere (tiny keeps the product off zero)
    ('disjoint', 0, [20, 21], [[10, 11]], ([0, 0, 0, 0], 2, 2)),
    # n-grams matched in DIFFERENT references all count: '10 11' from ref 1, '12 13' from ref 2,
    # '11 12' in neither
    ('matches_across_refs', 0, [10, 11, 12, 13], [[10, 11, 30, 31], [40, 12, 13, 41]], ([4, 2, 0, 0], 4, 4)),
    # use_eos = 1 keeps <end> as the word '0': three words on both sides
    ('eos_kept', 1, [10, 11, 0], [[10, 11, 0]], ([3, 2, 1, 0], 3, 3)),
    # use_eos = 0 drops it, <start> is skipped wherever it stands (utils.py:135-152)
    ('eos_dropped_start_skipped', 0, [1, 10, 1, 11, 0], [[1, 10, 11, 12, 0]], ([2, 1, 0, 0], 2, 3)),
    # longer than the reference: no penalty, unmatched tail
    ('longer_than_ref', 0, [10, 11, 12, 13, 14, 15], [[10, 11, 12]], ([3, 2, 1, 0], 6, 3)),
]
BLEU_ZERO = [   # the brevity penalty of an empty candidate underflows to exactly 0
    ('empty_candidate', 0, [], [[10, 11, 12]]),
    ('empty_candidate_empty_ref', 0, [], [[]]),
]

B2 = 1.2 ** 2


def rouge_closed(p, r):
    return ((1 + B2) * p * r) / float(r + B2 * p) if p != 0 and r != 0 else 0.0


# name, use_eos, candidate row, reference rows, (best precision, best recall) by hand
ROUGE_CASES = [
    ('identical', 0, [10, 11, 12], [[10, 11, 12]], (1.0, 1.0)),
    # LCS('10 11 12 13', '10 12 14') = '10 12'
    ('subsequence_not_substring', 0, [10, 11, 12, 13], [[10, 12, 14]], (2 / 4, 2 / 3)),
    # the two maxima come from DIFFERENT references: P = 4/4 from the long one, R = 2/2 from the short one
    ('maxima_from_different_refs', 0, [10, 11, 12, 13], [[10, 11, 12, 13, 14, 15, 16, 17], [10, 11]], (1.0, 1.0)),
    # order matters: reversed sentence shares a subsequence of length 1 only
    ('reversed', 0, [10, 11, 12], [[12, 11, 10]], (1 / 3, 1 / 3)),
    # repeated tokens: LCS('10 10 10', '10 10') = 2
    ('repeats', 0, [10, 10, 10], [[10, 10]], (2 / 3, 1.0)),
    ('disjoint', 0, [20, 21], [[10, 11]], (0.0, 0.0)),
    # split(" ") turns the empty string into ONE empty token: length 1, matching nothing here
    ('empty_candidate', 0, [], [[10, 11]], (0.0, 0.0)),
    # ... but it matches the empty token of an empty reference: LCS 1 of 1
    ('empty_vs_empty', 0, [], [[]], (1.0, 1.0)),
    # the empty reference gives R = 0/1, the other one P = 1/2 and R = 1/3
    ('empty_ref_among_refs', 0, [10, 20], [[], [10, 11, 12]], (1 / 2, 1 / 3)),
    # use_eos = 1: '10 0' vs '10 11 0' -> LCS 2
    ('eos_kept', 1, [10, 0], [[10, 11, 0]], (1.0, 2 / 3)),
    # interleaved: LCS('10 20 11 21 12', '10 11 12 13') = '10 11 12'
    ('interleaved', 0, [10, 20, 11, 21, 12], [[10, 11, 12, 13]], (3 / 5, 3 / 4)),
    # a 32-token reference fills the machine word of the bit-vector recurrence
    ('full_word', 0, list(range(100, 132)), [list(range(100, 132))], (1.0, 1.0)),
    ('full_word_shifted', 0, list(range(101, 133)), [list(range(100, 132))], (31 / 32, 31 / 32)),
]


def _canon(ids, use_eos):
    return ro.row_to_str(list(ids) + [0], use_eos)


@pytest.mark.parametrize('name,use_eos,hyp,refs,stats', BLEU_CASES, ids=[c[0] for c in BLEU_CASES])
def test_oracle_bleu_closed_forms(name, use_eos, hyp, refs, stats):
    _, lists = bo.Bleu().compute_score({0: [_canon(r, use_eos) for r in refs]}, {0: [_canon(hyp, use_eos)]})
    want = bleu_closed(*stats)
    for k in range(4):
        assert lists[k][0] == pytest.approx(want[k], rel=1e-12, abs=1e-300), (name, k)


@pytest.mark.parametrize('name,use_eos,hyp,refs', BLEU_ZERO, ids=[c[0] for c in BLEU_ZERO])
def test_oracle_bleu_of_an_empty_candidate_is_zero(name, use_eos, hyp, refs):
    _, lists = bo.Bleu().compute_score({0: [_canon(r, use_eos) for r in refs]}, {0: [_canon(hyp, use_eos)]})
    assert [l[0] for l in lists] == [0.0] * 4


@pytest.mark.parametrize('name,use_eos,hyp,refs,pr', ROUGE_CASES, ids=[c[0] for c in ROUGE_CASES])
def test_oracle_rouge_closed_forms(name, use_eos, hyp, refs, pr):
    _, got = bo.Rouge().compute_score({0: [_canon(r, use_eos) for r in refs]}, {0: [_canon(hyp, use_eos)]})
    assert got[0] == rouge_closed(*pr)


def test_oracle_corpus_bleu_adds_the_statistics():
    """The corpus score is built from the SUMS of the sentence statistics, not from the sentence scores."""
    gts = {0: ['a b c d e'], 1: ['a b'], 2: ['x y z w']}
    res = {0: ['a b c d e'], 1: ['a c'], 2: ['x y z']}
    # correct: [5,4,3,2] + [1,0,0,0] + [3,2,1,0]; guess: [5,4,3,2] + [2,1,0,0] + [3,2,1,0]; lengths 10 vs 11
    correct, guess = [9, 6, 4, 2], [10, 7, 4, 2]
    ratio = (10 + TINY) / (11 + SMALL)
    want, prod = [], 1.0
    for k in range(4):
        prod *= (correct[k] + TINY) / (guess[k] + SMALL)
        want.append(prod ** (1.0 / (k + 1)) * math.exp(1 - 1 / ratio))
    got, _ = bo.Bleu().compute_score(gts, res)
    assert got == pytest.approx(want, rel=1e-12)


def test_lcs_table_against_brute_force():
    rng = np.random.default_rng(5)
    for _ in range(200):
        a = rng.integers(0, 4, rng.integers(0, 7)).tolist()
        b = rng.integers(0, 4, rng.integers(0, 7)).tolist()
        best = 0
        for mask in range(1 << len(a)):                       # every subsequence of a
            sub = [a[i] for i in range(len(a)) if mask >> i & 1]
            it = iter(b)
            if all(any(x == y for y in it) for x in sub):
                best = max(best, len(sub))
        assert bo.lcs_length(a, b) == best


# ---------------------------------------------------------------------------
# fixtures produced by the reference's utils.py with these scorers
# ---------------------------------------------------------------------------
@pytest.fixture(scope='module')
def metric_cases():
    return np.load(os.path.join(GOLDEN, 'metric_reward_cases.npz'))


def _split(rows, counts):
    out, at = [], 0
    for c in counts:
        out.append(rows[at:at + c].astype(np.int64))
        at += c
    return out


CST_TAGS = (('gt_k7', dict(scb_captions=7, scb_baseline=1, use_mixer=1)), ('star_k20', dict(scb_captions=20, scb_baseline=2)),
            ('star_k3', dict(scb_captions=3, scb_baseline=2)), ('none', dict(scb_captions=0)))


def _check_fixtures(z, make, get_cst, get_sc, to_rows, to_gts, unwrap, rtol):
    for name in ('bleu', 'rouge'):
        scorer = make(name)
        gts = to_gts([z['sc/refs'][v] for v in range(5)])
        for ue in (0, 1):
            rew, m, g = get_sc(to_rows(z['sc/sampled']), to_rows(z['sc/greedy']), gts, scorer, expand_feat=1,
                               seq_per_img=20, use_eos=ue)
            np.testing.assert_allclose(unwrap(rew), z['sc/%s/eos%d/rewards' % (name, ue)], rtol=rtol, atol=rtol)
            assert float(m) == pytest.approx(float(z['sc/%s/eos%d/m_score' % (name, ue)]), rel=1e-9)
            assert float(g) == pytest.approx(float(z['sc/%s/eos%d/b_score' % (name, ue)]), rel=1e-9)
        gts = to_gts([z['cst/refs'][v] for v in range(6)])
        for tag, kw in CST_TAGS:
            if kw.get('scb_baseline') == 1:
                kw = dict(kw, bcmrscores=z['cst/bcmrscores'])
            rew, m, b = get_cst(to_rows(z['cst/sampled']), gts, scorer, expand_feat=1, seq_per_img=20, use_eos=1, **kw)
            np.testing.assert_allclose(unwrap(rew), z['cst/%s/%s/rewards' % (name, tag)], rtol=rtol, atol=rtol)
            assert float(m) == pytest.approx(float(z['cst/%s/%s/m_score' % (name, tag)]), rel=1e-9)
            assert float(b) == pytest.approx(float(z['cst/%s/%s/b_score' % (name, tag)]), rel=1e-9, abs=1e-300)
        gts = to_gts(_split(z['ragged/refs'], z['ragged/ref_counts']))
        rew, m, b = get_cst(to_rows(z['ragged/sampled']), gts, scorer, expand_feat=0, seq_per_img=3, scb_captions=2,
                            scb_baseline=2, use_eos=0)
        np.testing.assert_allclose(unwrap(rew), z['ragged/%s/rewards' % name], rtol=rtol, atol=rtol)
        assert float(m) == pytest.approx(float(z['ragged/%s/m_score' % name]), rel=1e-9)
        assert float(b) == pytest.approx(float(z['ragged/%s/b_score' % name]), rel=1e-9)
        gts = to_gts(_split(z['edge/refs'], z['edge/ref_counts']))
        for ue in (0, 1):
            rew, m, _ = get_cst(to_rows(z['edge/sampled']), gts, scorer, expand_feat=1, seq_per_img=6, scb_captions=0,
                                use_eos=ue)
            np.testing.assert_allclose(unwrap(rew), z['edge/%s/eos%d/rewards' % (name, ue)], rtol=rtol, atol=rtol)
            assert float(m) == pytest.approx(float(z['edge/%s/eos%d/m_score' % (name, ue)]), rel=1e-9)


def test_oracle_rewards_match_the_reference_run(metric_cases):
    _check_fixtures(metric_cases, lambda n: bo.Bleu() if n == 'bleu' else bo.Rouge(), ro.get_cst_reward,
                    ro.get_self_critical_reward, lambda a: a.astype(np.int64),
                    lambda gts: [np.asarray(g).astype(np.int64) for g in gts], lambda r: r[:, 0], 1e-12)


# ---------------------------------------------------------------------------
# the CUDA path
# ---------------------------------------------------------------------------
def _rows(lists, width):
    out = np.zeros((len(lists), width), dtype=np.int32)
    for i, ids in enumerate(lists):
        out[i, :len(ids)] = ids
    return out


@pytest.fixture(scope='module')
def gpu():
    import torch
    import cst_captioning_b200 as cst
    return cst, torch, torch.device('cuda:0')


@pytest.mark.gpu
def test_cuda_bleu_known_answers(gpu):
    cst, torch, dev = gpu
    scorer = cst.Bleu(device=dev)
    for ue in (0, 1):
        cases = [c for c in BLEU_CASES if c[1] == ue] + [c + (None,) for c in BLEU_ZERO if c[1] == ue]
        cand = torch.from_numpy(_rows([c[2] for c in cases], 8)).to(dev)
        offsets = np.concatenate([[0], np.cumsum([len(c[3]) for c in cases])])
        refs = cst.RefSet(torch.from_numpy(_rows([r for c in cases for r in c[3]], 8)).to(dev), offsets=offsets)
        got, comps = scorer.bleu_rows(cand, refs, 1, use_eos=ue, comps=True)
        got, comps = got.cpu().numpy(), comps.cpu().numpy()
        for i, c in enumerate(cases):
            if c[4] is None:
                assert got[:, i].tolist() == [0.0] * 4, c[0]
                continue
            correct, testlen, reflen = c[4]
            assert comps[i].tolist() == [testlen, reflen] + correct, c[0]        # the integer half: exact
            np.testing.assert_allclose(got[:, i], bleu_closed(*c[4]), rtol=1e-12, atol=1e-300, err_msg=c[0])
            _, lists = scorer.compute_score({3: [_canon(r, ue) for r in c[3]]}, {3: [_canon(c[2], ue)]})
            np.testing.assert_allclose([l[0] for l in lists], bleu_closed(*c[4]), rtol=1e-12, atol=1e-300, err_msg=c[0])
        got64 = scorer.bleu_rows(cand.long(), cst.RefSet(refs.rows.long(), offsets=offsets), 1, use_eos=ue)
        np.testing.assert_array_equal(got64.cpu().numpy(), got)


@pytest.mark.gpu
def test_cuda_rouge_known_answers_bit_exact(gpu):
    cst, torch, dev = gpu
    scorer = cst.Rouge(device=dev)
    for ue in (0, 1):
        cases = [c for c in ROUGE_CASES if c[1] == ue]
        cand = torch.from_numpy(_rows([c[2] for c in cases], 32)).to(dev)
        offsets = np.concatenate([[0], np.cumsum([len(c[3]) for c in cases])])
        refs = cst.RefSet(torch.from_numpy(_rows([r for c in cases for r in c[3]], 32)).to(dev), offsets=offsets)
        got = scorer.score_rows(cand, refs, 1, use_eos=ue).cpu().numpy()
        for c, s in zip(cases, got):
            assert s == rouge_closed(*c[4]), c[0]
            _, arr = scorer.compute_score({3: [_canon(r, ue) for r in c[3]]}, {3: [_canon(c[2], ue)]})
            assert arr[0] == rouge_closed(*c[4]), c[0]


@pytest.mark.gpu
def test_cuda_rewards_match_the_reference_run(gpu, metric_cases):
    cst, torch, dev = gpu
    _check_fixtures(metric_cases, lambda n: cst.Bleu(device=dev) if n == 'bleu' else cst.Rouge(device=dev),
                    cst.get_cst_reward, cst.get_self_critical_reward, lambda a: torch.from_numpy(a).to(dev),
                    lambda gts: [np.asarray(g) for g in gts], lambda r: r[:, 0].double().cpu().numpy(), 1e-6)


def _random_batch(rng, n_videos, s, n_refs, cw, rw, vocab, long_frac, empty_frac, ragged):
    counts = rng.integers(1, n_refs + 1, n_videos) if ragged else np.full(n_videos, n_refs)
    refs = rng.integers(2, vocab, (int(counts.sum()), rw)).astype(np.int32)
    cand = rng.integers(2, vocab, (n_videos * s, cw)).astype(np.int32)
    for arr, w in ((refs, rw), (cand, cw)):
        n = arr.shape[0]
        ln = rng.integers(1, min(12, w) + 1, n)
        if w > 16:
            ln = np.where(rng.random(n) < long_frac, rng.integers(17, w + 1, n), ln)
        ln = np.where(rng.random(n) < empty_frac, 0, ln)
        arr[np.arange(w)[None, :] >= ln[:, None]] = 0
    return cand, refs, counts


SHAPES = [  # videos, cands/video, refs, cand width, ref width, vocab, long rows, empty rows, ragged refs
    (40, 20, 20, 28, 30, 60, 0.05, 0.02, False),      # the training shape, small vocabulary: many matches
    (7, 45, 5, 16, 16, 12, 0.0, 0.05, False),         # more candidates than one batch of 32 rows
    (9, 3, 70, 30, 32, 30, 0.3, 0.02, True),          # ragged reference counts beyond 32: several stages
    (5, 33, 40, 32, 32, 6, 1.0, 0.0, False),          # every row long: batches cut by the instance budget
    (30, 1, 1, 9, 7, 8, 0.0, 0.2, False),             # one candidate, one reference; many empty rows
]


@pytest.mark.gpu
@pytest.mark.parametrize('shape', SHAPES, ids=[str(i) for i in range(len(SHAPES))])
def test_cuda_scores_against_the_oracle_on_random_batches(gpu, shape):
    cst, torch, dev = gpu
    n_videos, s, n_refs, cw, rw, vocab, long_frac, empty_frac, ragged = shape
    rng = np.random.default_rng(1000 + n_videos)
    cand, refs, counts = _random_batch(rng, n_videos, s, n_refs, cw, rw, vocab, long_frac, empty_frac, ragged)
    greedy = np.roll(cand, 3, axis=0)
    gts = _split(refs, counts)
    offsets = np.concatenate([[0], np.cumsum(counts)])
    refset = cst.RefSet(torch.from_numpy(refs).to(dev), offsets=offsets)
    c, g = torch.from_numpy(cand).to(dev), torch.from_numpy(greedy).to(dev)
    for ue in (0, 1):
        both = list(cand.astype(np.int64)) + list(greedy.astype(np.int64))
        want_r = ro.score_rows(both, gts, bo.Rouge(), 1, s, ue, n_cand=len(cand))
        got_r = cst.Rouge(device=dev).score_rows(c, refset, s, use_eos=ue, greedy=g).cpu().numpy()
        np.testing.assert_array_equal(got_r, want_r)                                # bit-exact
        hyp, rs = ro._strings_for(both, gts, ue)
        _, lists = bo.Bleu().compute_score({i: rs[(i % len(cand)) // s] for i in range(len(both))},
                                           {i: hyp[i] for i in range(len(both))})
        got_b = cst.Bleu(device=dev).bleu_rows(c, refset, s, use_eos=ue, greedy=g).cpu().numpy()
        np.testing.assert_allclose(got_b, np.array(lists), rtol=1e-12, atol=1e-300)
    # other row dtypes give the same bits
    for dt in (torch.int64, torch.int16):
        rs2 = cst.RefSet(refset.rows.to(dt), offsets=offsets)
        np.testing.assert_array_equal(cst.Rouge(device=dev).score_rows(c.to(dt), rs2, s, use_eos=1).cpu().numpy(),
                                      got_r[:len(cand)])
        np.testing.assert_array_equal(cst.Bleu(device=dev).bleu_rows(c.to(dt), rs2, s, use_eos=1).cpu().numpy(),
                                      got_b[:, :len(cand)])


@pytest.mark.gpu
def test_cuda_string_protocol_on_caption_words(gpu):
    """compute_score(gts, res) on raw words, corpus BLEU included; ROUGE's split(" ") keeps empty tokens."""
    cst, torch, dev = gpu
    gts = {'v1': ['a man is playing a guitar', 'someone plays the guitar', 'a man plays guitar on stage'],
           'v2': ['a cat sits on the mat', 'the cat is on a mat'],
           'v3': ['two  spaces here', ''],
           'v4': ['x']}
    res = {'v1': ['a man plays a guitar'], 'v2': ['a dog sits on the mat mat'], 'v3': [''], 'v4': ['x  x']}
    want_c, want_l = bo.Bleu().compute_score(gts, res)
    got_c, got_l = cst.Bleu(device=dev).compute_score(gts, res)
    np.testing.assert_allclose(got_c, want_c, rtol=1e-12)
    np.testing.assert_allclose(np.array(got_l), np.array(want_l), rtol=1e-12, atol=1e-300)
    want_m, want_s = bo.Rouge().compute_score(gts, res)
    got_m, got_s = cst.Rouge(device=dev).compute_score(gts, res)
    np.testing.assert_array_equal(got_s, want_s)
    assert got_m == want_m


@pytest.mark.gpu
def test_cuda_metric_scorers_refuse_what_they_cannot_do(gpu):
    cst, torch, dev = gpu
    with pytest.raises(ValueError):
        cst.Bleu(n=3, device=dev)
    rows = torch.zeros((2, 4, 8), dtype=torch.int32, device=dev)
    table = cst.DfTable.from_rows(rows, use_eos=1)
    cooked = cst.CookedRefs.cook(rows, cst.CiderD(df=table), use_eos=1)
    with pytest.raises(TypeError):
        cst.get_cst_reward(torch.zeros((8, 8), dtype=torch.int32, device=dev), cooked, cst.Rouge(device=dev),
                           expand_feat=1, seq_per_img=4, scb_captions=0)
    with pytest.raises(TypeError):
        cst.get_cst_reward(torch.zeros((8, 8), dtype=torch.int32, device=dev), rows, object(), expand_feat=1,
                           seq_per_img=4)
    # ids outside [0, 65534] end the caption and are counted, like in cst_reward (stats[4])
    cand = torch.tensor([[5, 6, 70000, 7], [5, 6, 0, 0]], dtype=torch.int64, device=dev)
    refs2 = torch.tensor([[[5, 6, 0, 0]]], dtype=torch.int64, device=dev)
    for sc in (cst.Bleu(device=dev), cst.Rouge(device=dev)):
        rew, m, b, st = cst.get_cst_reward(cand, refs2, sc, expand_feat=1, seq_per_img=2, scb_captions=0,
                                           return_stats=True)
        assert float(st[4]) == 1.0 and float(rew[0, 0]) == float(rew[1, 0])
    # an empty shard is nothing to score
    rew, m, b = cst.get_cst_reward(torch.zeros((0, 8), dtype=torch.int32, device=dev), rows[:0], cst.Bleu(device=dev),
                                   expand_feat=1, seq_per_img=4, scb_captions=0)
    assert rew.shape == (0, 8)


@pytest.mark.gpu
def test_cuda_host_pipeline_takes_these_scorers(gpu):
    """Pinned host rows through HostRewardPipeline (upload overlapped with scoring) = one launch on device rows."""
    cst, torch, dev = gpu
    from cst_captioning_b200 import synth
    b = synth.make_batch(300, 20, 20, seed=77)
    h_c, h_r = torch.from_numpy(b['sampled']).pin_memory(), torch.from_numpy(b['refs']).pin_memory()
    for sc in (cst.Bleu(device=dev), cst.Rouge(device=dev)):
        want, wm, wb = cst.get_cst_reward(h_c.to(dev), h_r.to(dev), sc, expand_feat=1, seq_per_img=20, scb_captions=5,
                                          scb_baseline=2, use_eos=1)
        pipe = cst.HostRewardPipeline(sc, seq_per_img=20, scb_captions=5, scb_baseline=2, use_eos=1, slices=7)
        out, stats = pipe(h_c, h_r)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(out.numpy(), want[:, 0].cpu().numpy())
        assert float(stats[5]) == pytest.approx(float(wm), rel=1e-12)
        assert float(stats[6]) == pytest.approx(float(wb), rel=1e-12)


@pytest.mark.gpu
def test_cuda_consensus_columns_bleu_rouge(gpu, tmp_path):
    """The Bleu_4 and ROUGE_L columns of compute_scores.py:75-103 (caption i against the others) and the
    pickle the data loader reads them from."""
    import pickle
    cst, torch, dev = gpu
    from cst_captioning_b200 import consensus, synth
    rng = np.random.default_rng(31)
    corpus, _ = synth.make_refs(rng, 60, 7, vocab=50)
    refs = torch.from_numpy(corpus).to(dev)
    got = consensus.evalscores(refs)
    assert set(got) == {'Bleu_4', 'ROUGE_L', 'CIDEr'}
    strings = [[ro.row_to_str(corpus[v, r], 0) for r in range(7)] for v in range(60)]
    for i in range(7):
        preds = {v: [strings[v][i]] for v in range(60)}
        gts = {v: strings[v][:i] + strings[v][i + 1:] for v in range(60)}
        _, lists = bo.Bleu().compute_score(gts, preds)
        np.testing.assert_allclose(got['Bleu_4'][:, i].cpu().numpy(), np.array(lists[-1]), rtol=1e-12, atol=1e-300)
        _, want = bo.Rouge().compute_score(gts, preds)
        np.testing.assert_array_equal(got['ROUGE_L'][:, i].cpu().numpy(), want)
    host = consensus.write_evalscores_pkl(str(tmp_path / 'evalscores.pkl'), got)
    back = pickle.load(open(tmp_path / 'evalscores.pkl', 'rb'))
    assert set(back) == set(host) and back['ROUGE_L'].shape == (60, 7) and back['CIDEr'].dtype == np.float64
    with pytest.raises(ValueError):
        consensus.consensus_scores(refs, metric='METEOR')
