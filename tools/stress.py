"""Stress k_score2 (and k_bleu / k_rouge) over many random shapes (ragged lengths, long rows, more references than one table
takes, candidate counts beyond one batch): repeated launches must agree bit for bit (determinism),
the fused launch must agree with a per-candidate rescoring, and staged references with raw rows.
Run under `timeout`: a hang would show as a timeout, not as a wrong number."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as g
g.build()
import cst_captioning_b200 as cst
from cst_captioning_b200 import synth
from cst_captioning_b200.reward import RefSet

dev = torch.device('cuda:0')
rng = np.random.default_rng(99)
corpus, _ = synth.make_refs(rng, 500, 20, vocab=300)
scorer = cst.CiderD(df=cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1))
n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 150
metric_scorers = [cst.Bleu(device=dev), cst.Rouge(device=dev)]
t0 = time.time()
for case in range(n_cases):
    n_videos = int(rng.integers(1, 3000))
    s = int(rng.choice([1, 2, 3, 7, 20, 20, 20, 33, 41]))
    n_refs = int(rng.choice([1, 2, 5, 19, 20, 20, 21, 32, 33, 47, 70]))
    long_frac = float(rng.choice([0.0, 0.04, 0.3, 1.0]))
    cw, rw = int(rng.choice([8, 17, 28, 32])), int(rng.choice([9, 18, 30, 32]))
    refs = rng.integers(3, 300, (n_videos, n_refs, rw)).astype(np.int32)
    cand = rng.integers(3, 300, (n_videos * s, cw)).astype(np.int32)
    # lengths: short rows (2..14 words) or long ones (17..width)
    for arr, w in ((refs.reshape(-1, rw), rw), (cand, cw)):
        n = arr.shape[0]
        ln = rng.integers(2, min(15, w) + 1, n)
        is_long = rng.random(n) < long_frac
        if w > 17:
            ln = np.where(is_long, rng.integers(17, w + 1, n), ln)
        arr[np.arange(w)[None, :] >= ln[:, None]] = 0
    refs[:, :, 0] = 1
    c = torch.from_numpy(cand).to(dev)
    r = torch.from_numpy(refs).to(dev)
    k = int(rng.integers(1, s + 1))
    ue = int(rng.integers(0, 2))
    outs = []
    for rep in range(3):
        rew, m, b, st = cst.get_cst_reward(c, r, scorer, expand_feat=1, seq_per_img=s, scb_captions=k,
                                           scb_baseline=2, use_eos=ue, return_stats=True)
        outs.append((rew[:, 0].clone(), st.clone()))
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(o[0], outs[0][0]) and torch.equal(o[1], outs[0][1]), 'non-deterministic result in case %d' % case
    # scores of the fused launch == scores of one candidate per "video" (different CTA shapes)
    a = scorer.score_rows(c, RefSet.coerce(r, dev), s, use_eos=ue)
    per = RefSet.coerce(r.repeat_interleave(s, dim=0)[:min(64, n_videos * s)], dev)
    bsc = scorer.score_rows(c[:per.n_videos], per, 1, use_eos=ue)
    assert torch.equal(a[:per.n_videos], bsc), 'fused vs per-candidate mismatch in case %d' % case
    if n_refs <= 32:                 # references staged once give the bits of the raw rows
        cooked = cst.CookedRefs.cook(r, scorer, use_eos=ue)
        ids = torch.from_numpy(rng.permutation(n_videos)).to(dev)
        rew2, _, _, st2 = cst.get_cst_reward(c.reshape(n_videos, s, cw)[ids].reshape(-1, cw), cooked.batch(ids), scorer,
                                             expand_feat=1, seq_per_img=s, scb_captions=k, scb_baseline=2,
                                             use_eos=ue, return_stats=True)
        want = outs[0][0].reshape(n_videos, s)[ids].reshape(-1)
        # a video whose references exceed one table (1,100 n-gram instances) cannot be staged: its image is
        # invalid, counted at cook time, and its rewards are NaN -- every other video must match bit for bit
        nan_videos = torch.isnan(rew2[:, 0]).reshape(n_videos, s).all(1)
        assert int(nan_videos.sum()) == int(cooked.bad_rows.item()), 'invalid images miscounted in case %d' % case
        ok = ~nan_videos.repeat_interleave(s)
        assert torch.equal(rew2[:, 0][ok], want[ok]), 'staged vs raw mismatch in case %d' % case
        assert not bool(torch.isnan(want).any())
        del cooked
    g2 = torch.roll(c, 1, 0)
    cst.get_self_critical_reward(c, g2, r, scorer, expand_feat=1, seq_per_img=s, use_eos=ue)
    # the BLEU-4 / ROUGE-L kernels: repeated launches agree bit for bit, and the batched launch (candidate
    # batches, reference stages) gives the scores of one candidate per "video" (a single short stage each)
    rs = RefSet.coerce(r, dev)
    for sc in metric_scorers:
        a1 = sc.score_rows(c, rs, s, use_eos=ue, greedy=g2)
        a2 = sc.score_rows(c, rs, s, use_eos=ue, greedy=g2)
        assert torch.equal(a1, a2), '%s: non-deterministic result in case %d' % (type(sc).__name__, case)
        b1 = sc.score_rows(c[:per.n_videos], per, 1, use_eos=ue)
        assert torch.equal(a1[:per.n_videos], b1), '%s: batched vs per-candidate mismatch in case %d' % (type(sc).__name__, case)
        assert bool(torch.isfinite(a1).all()) and float(a1.min()) >= 0.0 and float(a1.max()) <= 1.0 + 1e-9
torch.cuda.synchronize()
print('stress ok: %d cases in %.1f s' % (n_cases, time.time() - t0))
