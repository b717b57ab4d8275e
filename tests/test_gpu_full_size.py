"""The BASELINE.json configs at their STATED sizes against the CPU oracle (``-m gpu``).

The fixtures of ``tests/golden`` are reduced-size cases; here the synthetic corpus of SURVEY.md §8d
is scored at full size on both sides.  Float64 scores are compared directly (relative 1e-9: the
north_star bar is 1e-5), rewards through the float32 column the kernel writes.  The oracle needs
about two minutes for the whole file.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DF_VIDEOS = 6513


@pytest.fixture(scope='module')
def world():
    import __graft_entry__ as graft
    graft.build()
    import cst_captioning_b200 as cst
    from cst_captioning_b200 import synth
    from oracle import ciderd_oracle as co
    dev = torch.device('cuda:0')
    rng = np.random.default_rng(synth.SEED)
    corpus, n_words = synth.make_refs(rng, DF_VIDEOS, 20)
    table = cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1)
    df, n_doc = co.build_df_from_strings(synth.df_corpus_strings(corpus, n_words))
    assert len(table) == len(df) and n_doc == DF_VIDEOS
    return cst, dev, cst.CiderD(df=table), co.CiderD(df=(df, n_doc)), corpus


def _close64(got, want):
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-13)


def test_config3_self_critical_1280_plus_1280(world):
    """CST_MS_Greedy: batch 64 x 20 sampled + 20 greedy captions, reward = sample - greedy."""
    cst, dev, scorer, oracle, _ = world
    from cst_captioning_b200 import synth
    from oracle import reward_oracle as ro
    b = synth.make_batch(64, 20, 20, seed=12, greedy=True)
    gts = [b['refs'][v].astype(np.int64) for v in range(64)]
    want, wm, wg = ro.get_self_critical_reward(b['sampled'].astype(np.int64), b['greedy'].astype(np.int64), gts, oracle,
                                               expand_feat=1, seq_per_img=20, use_eos=1)
    c, g, r = (torch.from_numpy(b[k]).to(dev) for k in ('sampled', 'greedy', 'refs'))
    rew, m, gs = cst.get_self_critical_reward(c.long(), g.long(), r.long(), scorer, expand_feat=1, seq_per_img=20, use_eos=1)
    np.testing.assert_allclose(rew[:, 0].double().cpu().numpy(), want[:, 0].astype(np.float32), rtol=1e-5, atol=1e-6)
    _close64(float(m), wm)
    _close64(float(gs), wg)
    from cst_captioning_b200.reward import RefSet
    scores = scorer.score_rows(c, RefSet.coerce(r, dev), 20, use_eos=1, greedy=g).cpu().numpy()
    want_s = ro.score_rows(list(b['sampled'].astype(np.int64)) + list(b['greedy'].astype(np.int64)), gts, oracle, 1, 20, 1,
                           n_cand=1280)
    _close64(scores, want_s)


def test_config4_scb_gt_batch_512(world):
    """CST_MS_SCB: SCB_CAPTIONS = 20, baseline from GT consensus scores, batch 512."""
    cst, dev, scorer, oracle, _ = world
    from cst_captioning_b200 import synth
    from oracle import reward_oracle as ro
    b = synth.make_batch(512, 20, 20, seed=13)
    bcmr = np.random.default_rng(4).random((512, 20)) * 3.0
    gts = [b['refs'][v].astype(np.int64) for v in range(512)]
    want, wm, wb = ro.get_cst_reward(b['sampled'].astype(np.int64), gts, oracle, bcmrscores=bcmr, expand_feat=1,
                                     seq_per_img=20, scb_captions=20, scb_baseline=1, use_eos=1, use_mixer=1)
    rew, m, bs = cst.get_cst_reward(torch.from_numpy(b['sampled']).to(dev), torch.from_numpy(b['refs']).to(dev), scorer,
                                    bcmrscores=bcmr, expand_feat=1, seq_per_img=20, scb_captions=20, scb_baseline=1,
                                    use_eos=1, use_mixer=1)
    np.testing.assert_allclose(rew[:, 0].double().cpu().numpy(), want[:, 0].astype(np.float32), rtol=1e-5, atol=1e-6)
    _close64(float(m), wm)
    _close64(float(bs), wb)


def test_config2_consensus_scores_all_6513_videos_two_columns(world):
    """CST_GT_None / WXE weights: each GT caption scored against the other 19 with a corpus-mode df
    (compute_scores.py:75-103), all 6,513 videos x 20.  The device computes every column from ONE table
    plus leave-one-out counts; the oracle rebuilds the df per column like the reference, so two
    columns of it are affordable here -- and on a slice of videos all twenty are checked against a
    second device path that rebuilds the table per pass."""
    cst, dev, _, _, corpus = world
    from cst_captioning_b200.consensus import consensus_scores
    from oracle import ciderd_oracle as co
    from oracle import reward_oracle as ro
    refs = torch.from_numpy(corpus).to(dev)
    got = consensus_scores(refs).cpu().numpy()                       # (6513, 20)
    assert got.shape == (DF_VIDEOS, 20) and np.isfinite(got).all() and got.min() >= 0
    strings = [[ro.row_to_str(corpus[v, r], 0) for r in range(20)] for v in range(DF_VIDEOS)]
    vids = list(range(DF_VIDEOS))
    for i in (0, 13):
        preds = {v: [strings[v][i]] for v in vids}
        gts = {v: strings[v][:i] + strings[v][i + 1:] for v in vids}
        _, want = co.Cider().compute_score(gts, preds)
        _close64(got[:, i], want)
    # the per-pass-table path (what videos beyond one shared-memory table fall back to) agrees on a slice
    import cst_captioning_b200 as cstp
    sub = refs[:200].contiguous()
    fast = consensus_scores(sub).cpu().numpy()
    slow = np.zeros_like(fast)
    for i in range(20):
        others = torch.cat([sub[:, :i], sub[:, i + 1:]], dim=1).contiguous()
        t_i = cstp.DfTable.from_rows(others, use_eos=0)
        rs = cstp.RefSet(others.reshape(200 * 19, -1), per_video=19)
        slow[:, i] = cstp.CiderD(df=t_i).score_rows(sub[:, i].contiguous(), rs, 1, use_eos=0).cpu().numpy()
        t_i.close()
    _close64(fast, slow)


def test_config5_scb_star_4096_videos_properties_and_slice(world):
    """CST_MS_SCB*: 20 samples per video, leave-nothing-out baseline over the k = 20 lowest (all): the
    rewards of a video sum to zero; a slice of 64 videos is held to the oracle in float64."""
    cst, dev, scorer, oracle, _ = world
    from cst_captioning_b200 import synth
    from cst_captioning_b200.reward import RefSet
    from oracle import reward_oracle as ro
    b = synth.make_batch(4096, 20, 20, seed=1000)
    c, r = torch.from_numpy(b['sampled']).to(dev), torch.from_numpy(b['refs']).to(dev)
    rew, m, bs, st = cst.get_cst_reward(c, r, scorer, expand_feat=1, seq_per_img=20, scb_captions=20, scb_baseline=2,
                                        use_eos=1, return_stats=True)
    assert float(rew[:, 0].reshape(4096, 20).double().sum(1).abs().max()) < 1e-4
    assert float(m) == pytest.approx(float(bs), rel=1e-12) and float(st[4]) == 0.0
    lo, hi = 3000, 3064
    gts = [b['refs'][v].astype(np.int64) for v in range(lo, hi)]
    want = ro.score_rows(list(b['sampled'][lo * 20:hi * 20].astype(np.int64)), gts, oracle, 1, 20, 1)
    scores = scorer.score_rows(c, RefSet.coerce(r, dev), 20, use_eos=1).cpu().numpy()
    _close64(scores[lo * 20:hi * 20], want)


def test_other_scorers_4096_videos_properties_and_slice(world):
    """BLEU-4 / ROUGE-L as the reward scorer (train.py:119-124) on the config-5 shape: properties over all
    81,920 captions (scores in [0, 1]; a caption scored against references that contain it gets ROUGE-L 1
    and BLEU-1..3 of 1 up to the scorer's 1e-9 offsets; SCB* rewards of a video sum to zero), and a slice of
    64 videos held to the restated scorers."""
    cst, dev, _, _, _ = world
    from cst_captioning_b200 import synth
    from cst_captioning_b200.reward import RefSet
    from oracle import bleu_rouge_oracle as bo
    from oracle import reward_oracle as ro
    b = synth.make_batch(4096, 20, 20, seed=1000)
    c, r = torch.from_numpy(b['sampled']).to(dev), torch.from_numpy(b['refs']).to(dev)
    refset = RefSet.coerce(r, dev)
    bleu, rouge = cst.Bleu(device=dev), cst.Rouge(device=dev)
    lo, hi = 2000, 2064
    gts = [b['refs'][v].astype(np.int64) for v in range(lo, hi)]
    rows = list(b['sampled'][lo * 20:hi * 20].astype(np.int64))
    for sc, theirs, exact in ((bleu, bo.Bleu(), False), (rouge, bo.Rouge(), True)):
        s = sc.score_rows(c, refset, 20, use_eos=1)
        assert float(s.min()) >= 0.0 and float(s.max()) <= 1.0 and bool(torch.isfinite(s).all())
        want = ro.score_rows(rows, gts, theirs, 1, 20, 1)
        got = s[lo * 20:hi * 20].cpu().numpy()
        if exact:
            np.testing.assert_array_equal(got, want)
        else:
            np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-300)
        rew, m, bs = cst.get_cst_reward(c, r, sc, expand_feat=1, seq_per_img=20, scb_captions=20, scb_baseline=2, use_eos=1)
        assert float(rew[:, 0].reshape(4096, 20).double().sum(1).abs().max()) < 1e-5
        assert float(m) == pytest.approx(float(bs), rel=1e-12)
    # every reference scored against its own video's references
    own = r.reshape(4096 * 20, -1)
    assert bool((rouge.score_rows(own, refset, 20, use_eos=1) == 1.0).all())
    b123 = bleu.bleu_rows(own, refset, 20, use_eos=1)[:3]
    assert float((b123 - 1.0).abs().max()) < 1e-8
