"""cst_reward_noalloc -- the reward launch for hosts that cannot allocate.

Some hosts hand all of HBM to their own memory pool at start-up (a plugin's
cudaMalloc then fails) or forbid allocation inside the step loop because
cudaMalloc / cudaFree stall the device.  For them the library keeps the statistics
scratch in a fixed array of its own.  This test drives that entry point through the
C ABI and checks it against cst_reward (caller-provided scratch: bit for bit) and against
the CPU oracle on two calls in a row, with different sizes and baseline modes, so that
anything the first call leaves in the shared array would show up in the second; then the
path that runs two kernels over the array, before and after a call the library refuses.

The file name makes pytest collect it last: if it fails under ``-x`` every other
test has already run.  It passes alone in a fresh process as well as after the
other tests.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def env():
    import __graft_entry__ as graft
    graft.build()
    import cst_captioning_b200 as cst
    from cst_captioning_b200 import synth
    dev = torch.device('cuda:0')
    rng = np.random.default_rng(5)
    corpus, _ = synth.make_refs(rng, 300, 20)
    table = cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1)
    return cst, dev, table


def _call(fn, table, dev, batch, mode, k, with_scratch, cands_per_video=20, seq_per_img=20):
    """One launch through the C ABI -> (reward float32, stats float64, scores float64 or None) as numpy."""
    from cst_captioning_b200 import _lib
    cand = torch.from_numpy(batch['sampled']).to(dev)
    greedy = torch.from_numpy(batch['greedy']).to(dev) if mode == _lib.BASE_GREEDY else None
    refs = torch.from_numpy(batch['refs']).to(dev)
    n_entries, n_refs, ref_w = refs.shape
    n = cand.shape[0]
    assert n == n_entries * cands_per_video
    reward = torch.full((n,), float('nan'), dtype=torch.float32, device=dev)
    stats = torch.full((_lib.N_STATS,), float('nan'), dtype=torch.float64, device=dev)
    # baseline rows that are not videos (cands_per_video != seq_per_img) go through two kernels and
    # need the scores buffer: score, then baseline -- both share the scratch
    scores = None
    if cands_per_video != seq_per_img:
        scores = torch.full((n * (2 if greedy is not None else 1),), float('nan'), dtype=torch.float64, device=dev)
    args = [table.handle, cand.data_ptr(), greedy.data_ptr() if greedy is not None else None,
            _lib.DTYPE_I32, n, cand.shape[1], cand.stride(0),
            refs.data_ptr(), _lib.DTYPE_I32, ref_w, ref_w, None, n_refs, n_entries, cands_per_video, 1,
            mode, seq_per_img, k, None, reward.data_ptr(), scores.data_ptr() if scores is not None else None,
            stats.data_ptr()]
    if with_scratch:
        nbytes = int(_lib.lib.cst_reward_scratch_bytes(max(n_entries, n // seq_per_img)))
        scratch = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        args += [scratch.data_ptr(), nbytes]
    args.append(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(fn(*args))
    torch.cuda.synchronize()
    return reward.cpu().numpy(), stats.cpu().numpy(), None if scores is None else scores.cpu().numpy()


def _oracle(table, cst, batch, mode, k, expand_feat=1, seq_per_img=20):
    """The same call on the CPU oracle (test infrastructure): float64 rewards, m_score, b_score."""
    from cst_captioning_b200 import _lib
    from oracle import ciderd_oracle as co
    from oracle import reward_oracle as ro
    keys, df = table.export()
    scorer = co.CiderD(df=(dict(zip(cst.unpack_keys(keys), df.tolist())), table.n_documents))
    gts = [batch['refs'][v].astype(np.int64) for v in range(batch['refs'].shape[0])]
    if mode == _lib.BASE_GREEDY:
        rew, m, b = ro.get_self_critical_reward(batch['sampled'].astype(np.int64), batch['greedy'].astype(np.int64),
                                                gts, scorer, expand_feat=expand_feat, seq_per_img=seq_per_img, use_eos=1)
    else:
        rew, m, b = ro.get_cst_reward(batch['sampled'].astype(np.int64), gts, scorer, expand_feat=expand_feat,
                                      seq_per_img=seq_per_img, scb_captions=k, scb_baseline=2, use_eos=1)
    return rew[:, 0], m, b


def test_noalloc_matches_cst_reward_twice_in_a_row(env):
    cst, dev, table = env
    from cst_captioning_b200 import _lib, synth
    lib = _lib.lib
    first = synth.make_batch(96, 20, 20, seed=31, greedy=True)
    second = synth.make_batch(37, 20, 20, seed=32, greedy=True)
    plan = [(first, _lib.BASE_SCB_SAMPLE, 20), (second, _lib.BASE_GREEDY, 0)]
    for batch, mode, k in plan:                       # two calls in a row share the built-in array
        got_r, got_s, _ = _call(lib.cst_reward_noalloc, table, dev, batch, mode, k, with_scratch=False)
        want_r, want_s, _ = _call(lib.cst_reward, table, dev, batch, mode, k, with_scratch=True)
        assert np.isfinite(got_r).all() and np.isfinite(got_s).all()
        # same kernels, same order of operations: where the scratch lives must not change a bit ...
        np.testing.assert_array_equal(got_r, want_r)
        np.testing.assert_array_equal(got_s, want_s)
        assert got_s[4] == 0.0
        # ... and the numbers are the CPU oracle's, within the tolerance of the parity tests
        o_r, o_m, o_b = _oracle(table, cst, batch, mode, k)
        np.testing.assert_allclose(got_r, o_r.astype(np.float32), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(got_s[5], o_m, rtol=1e-9)
        np.testing.assert_allclose(got_s[6], o_b, rtol=1e-9)


def test_noalloc_two_kernel_path_and_after_a_refused_call(env):
    """One candidate per reference entry with baseline rows of 5 captions (expand_feat = 0,
    utils.py:271-272): the scoring kernel and the baseline kernel both take the built-in array.
    A call the library refuses in between must leave the array usable."""
    cst, dev, table = env
    from cst_captioning_b200 import _lib, synth
    lib = _lib.lib
    base = synth.make_batch(45, 1, 20, seed=33)                 # 45 entries x 1 caption, rows of 5
    for attempt in range(2):
        got_r, got_s, got_sc = _call(lib.cst_reward_noalloc, table, dev, base, _lib.BASE_SCB_SAMPLE, 3,
                                     with_scratch=False, cands_per_video=1, seq_per_img=5)
        want_r, want_s, want_sc = _call(lib.cst_reward, table, dev, base, _lib.BASE_SCB_SAMPLE, 3,
                                        with_scratch=True, cands_per_video=1, seq_per_img=5)
        np.testing.assert_array_equal(got_r, want_r)
        np.testing.assert_array_equal(got_s, want_s)
        np.testing.assert_array_equal(got_sc, want_sc)
        o_r, o_m, o_b = _oracle(table, cst, base, _lib.BASE_SCB_SAMPLE, 3, expand_feat=0, seq_per_img=5)
        np.testing.assert_allclose(got_r, o_r.astype(np.float32), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(got_s[5], o_m, rtol=1e-9)
        np.testing.assert_allclose(got_s[6], o_b, rtol=1e-9)
        if attempt == 0:
            # scb_k beyond the row length: refused before anything is launched
            rc = lib.cst_reward_noalloc(table.handle, C.c_void_p(256), None, 0, 45, 28, 28, C.c_void_p(256), 0, 30,
                                        30, None, 20, 45, 1, 1, _lib.BASE_SCB_SAMPLE, 5, 6, None, C.c_void_p(256),
                                        C.c_void_p(256), C.c_void_p(256), None)
            assert rc == _lib.E_ARG


def test_noalloc_refuses_batches_beyond_its_array(env):
    cst, dev, table = env
    from cst_captioning_b200 import _lib
    lib = _lib.lib
    too_many = 16384 + 1                               # CST_NOALLOC_MAX_GROUPS + 1
    rc = lib.cst_reward_noalloc(table.handle, C.c_void_p(256), None, 0, too_many * 20, 28, 28,
                                C.c_void_p(256), 0, 30, 30, None, 20, too_many, 20, 1,
                                _lib.BASE_NONE, 20, 0, None, C.c_void_p(256), None, C.c_void_p(256), None)
    assert rc == _lib.E_RANGE
    assert b'CST_NOALLOC_MAX_GROUPS' in lib.cst_last_error()
