"""cst_reward_noalloc -- the reward launch for hosts that cannot allocate.

Some hosts hand all of HBM to their own memory pool at start-up (a plugin's
cudaMalloc then fails) or forbid allocation inside the step loop because
cudaMalloc / cudaFree stall the device.  For them the library keeps the statistics
scratch in a fixed array of its own.  This test drives that entry point through the
C ABI and checks it against cst_reward (caller-provided scratch) on two calls in a
row, with different sizes and baseline modes, so that anything the first call leaves
in the shared array would show up in the second.

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


def _call(fn, table, dev, batch, mode, k, with_scratch):
    """One launch through the C ABI -> (reward float32, stats float64) as numpy."""
    from cst_captioning_b200 import _lib
    cand = torch.from_numpy(batch['sampled']).to(dev)
    greedy = torch.from_numpy(batch['greedy']).to(dev) if mode == _lib.BASE_GREEDY else None
    refs = torch.from_numpy(batch['refs']).to(dev)
    n_videos, n_refs, ref_w = refs.shape
    n = cand.shape[0]
    reward = torch.full((n,), float('nan'), dtype=torch.float32, device=dev)
    stats = torch.full((_lib.N_STATS,), float('nan'), dtype=torch.float64, device=dev)
    args = [table.handle, cand.data_ptr(), greedy.data_ptr() if greedy is not None else None,
            _lib.DTYPE_I32, n, cand.shape[1], cand.stride(0),
            refs.data_ptr(), _lib.DTYPE_I32, ref_w, ref_w, None, n_refs, n_videos, 20, 1,
            mode, 20, k, None, reward.data_ptr(), None, stats.data_ptr()]
    if with_scratch:
        nbytes = int(_lib.lib.cst_reward_scratch_bytes(n_videos))
        scratch = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        args += [scratch.data_ptr(), nbytes]
    args.append(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(fn(*args))
    torch.cuda.synchronize()
    return reward.cpu().numpy(), stats.cpu().numpy()


def test_noalloc_matches_cst_reward_twice_in_a_row(env):
    cst, dev, table = env
    from cst_captioning_b200 import _lib, synth
    lib = _lib.lib
    first = synth.make_batch(96, 20, 20, seed=31, greedy=True)
    second = synth.make_batch(37, 20, 20, seed=32, greedy=True)
    plan = [(first, _lib.BASE_SCB_SAMPLE, 20), (second, _lib.BASE_GREEDY, 0)]
    for batch, mode, k in plan:                       # two calls in a row share the built-in array
        got_r, got_s = _call(lib.cst_reward_noalloc, table, dev, batch, mode, k, with_scratch=False)
        want_r, want_s = _call(lib.cst_reward, table, dev, batch, mode, k, with_scratch=True)
        assert np.isfinite(got_r).all() and np.isfinite(got_s).all()
        # same kernels, same order of operations: the tolerance of the parity tests
        # (1e-5 relative) is far looser than what is observed (bit-identical)
        np.testing.assert_allclose(got_r, want_r, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(got_s, want_s, rtol=1e-5, atol=1e-9)
        assert got_s[4] == 0.0


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
