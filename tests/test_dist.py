"""Host-side logic of the multi-GPU path on CPU: world_size 2, gloo backend.

The CUDA kernels are not involved (no GPU here); each rank computes its slice's
statistics with the oracle, and the product's sharding + all-reduce code must
reproduce the single-process result of the reference's reward function."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import GOLDEN, ids_to_df_dict


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        import __graft_entry__ as graft
        graft.build()
        from cst_captioning_b200 import dist as cdist
        from oracle import ciderd_oracle as co
        from oracle import reward_oracle as ro
        g = np.load(os.path.join(GOLDEN, 'reward_cases.npz'))
        z = np.load(os.path.join(GOLDEN, 'df_v10k.npz'))
        scorer = co.CiderD(df=(ids_to_df_dict(z['ngram_ids'], z['df']), int(z['ref_len'])))
        refs, samp = g['cfg5/refs'], g['cfg5/sampled']
        part = cdist.shard_batch(samp, refs, 20, rank, world)
        lo, hi = part['videos']
        gts = [part['data_gts'][v].astype(np.int64) for v in range(hi - lo)]
        k = 3
        rew, m, b = ro.get_cst_reward(part['model_res'].astype(np.int64), gts, scorer, expand_feat=1,
                                      seq_per_img=20, scb_captions=k, scb_baseline=2, use_eos=1)
        n_loc = (hi - lo) * 20
        # what the kernel writes per rank: sums and counts (include/cst_reward.h, stats_out)
        stats = torch.tensor([m * n_loc, b * (hi - lo) * k, n_loc, (hi - lo) * k, 0.0], dtype=torch.float64)
        cdist.all_reduce_stats(stats)
        gm, gb = cdist.means(stats)
        np.save(os.path.join(out_dir, 'rew%d.npy' % rank), rew[:, 0])
        if rank == 0:
            np.save(os.path.join(out_dir, 'means.npy'), np.array([float(gm), float(gb), lo, hi]))
    finally:
        dist.destroy_process_group()


def test_shard_range_covers_everything():
    from cst_captioning_b200.dist import shard_range
    for n in (0, 1, 7, 64, 65536, 65537):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_two_rank_gloo_matches_single_process(tmp_path, golden_cases):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    g = golden_cases
    rew = np.concatenate([np.load(str(tmp_path / ('rew%d.npy' % r))) for r in range(world)])
    # rewards are per video: sharding must not change a single bit
    np.testing.assert_array_equal(rew, g['cfg5/k3/eos1/rewards'])
    gm, gb, lo, hi = np.load(str(tmp_path / 'means.npy'))
    assert (lo, hi) == (0, 3)
    np.testing.assert_allclose(gm, float(g['cfg5/k3/eos1/m_score']), rtol=1e-12)
    np.testing.assert_allclose(gb, float(g['cfg5/k3/eos1/b_score']), rtol=1e-12)
