"""The sharded reward on two real GPUs over NCCL against one GPU scoring the whole batch
(``-m gpu``; skipped on a box with a single device -- ``bench.py --gpus N`` makes the same check
at N = 2, 4, 8 and reports it as ``strong``)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    try:
        import cst_captioning_b200 as cst
        from cst_captioning_b200 import dist as cdist
        from cst_captioning_b200 import synth
        rng = np.random.default_rng(synth.SEED)
        corpus, _ = synth.make_refs(rng, 300, 20)
        scorer = cst.CiderD(df=cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1))
        b = synth.make_batch(1001, 20, 20, seed=77)            # an odd count: the shards differ in size
        cand, refs = torch.from_numpy(b['sampled']).to(dev), torch.from_numpy(b['refs']).to(dev)
        rew, m, bs, (lo, hi) = cdist.get_cst_reward_sharded(cand, refs, scorer, seq_per_img=20, scb_captions=7,
                                                            scb_baseline=2, use_eos=1)
        full = cst.get_cst_reward(cand, refs, scorer, expand_feat=1, seq_per_img=20, scb_captions=7, scb_baseline=2,
                                  use_eos=1)
        assert torch.equal(rew[:, 0], full[0][lo * 20:hi * 20, 0])           # rewards are per video: same bits
        np.save(os.path.join(out_dir, 'means%d.npy' % rank),
                np.array([float(m), float(bs), float(full[1]), float(full[2]), lo, hi]))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_two_gpu_nccl_matches_one_gpu(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    spans = []
    for r in range(world):
        m, b, m1, b1, lo, hi = np.load(str(tmp_path / ('means%d.npy' % r)))
        np.testing.assert_allclose([m, b], [m1, b1], rtol=1e-12)
        spans.append((int(lo), int(hi)))
    assert spans == [(0, 501), (501, 1001)]
