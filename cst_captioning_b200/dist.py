"""Data parallelism over videos.

The reference is single-process (no collective anywhere).  A video's rewards depend
only on its own candidates and references plus the global df table, and the SCB
baselines are per video (``utils.py:308-315``), so the batch shards by video with no
data-path exchange: rank g scores a contiguous slice, the df table is replicated, and
the only thing ranks share is the pair of logged means (``train.py:229-233``), i.e. an
all-reduce of the statistics sums the kernel writes.

One process per GPU (``torchrun``); NCCL on the GPUs, gloo in the CPU tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n_videos, rank, world):
    """Contiguous, balanced slice [lo, hi) of the videos for ``rank``."""
    if not 0 <= rank < world:
        raise ValueError('rank %d outside world of %d' % (rank, world))
    base, extra = divmod(int(n_videos), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(model_res, data_gts, seq_per_img, rank, world, bcmrscores=None, greedy_res=None):
    """Slice a global batch laid out as ``expand_feat=1`` wants it (candidate i belongs
    to video ``i // seq_per_img``, ``utils.py:266-270``) -> dict of this rank's views."""
    n_videos = len(data_gts) if not torch.is_tensor(data_gts) else data_gts.shape[0]
    if model_res.shape[0] != n_videos * seq_per_img:
        raise ValueError('model_res must hold seq_per_img rows per video')
    lo, hi = shard_range(n_videos, rank, world)
    out = {'videos': (lo, hi), 'model_res': model_res[lo * seq_per_img:hi * seq_per_img],
           'data_gts': data_gts[lo:hi]}
    if bcmrscores is not None:
        out['bcmrscores'] = bcmrscores[lo:hi]
    if greedy_res is not None:
        out['greedy_res'] = greedy_res[lo * seq_per_img:hi * seq_per_img]
    return out


def all_reduce_stats(stats, group=None):
    """Sum the statistics ``[sum_m, sum_b, n_m, n_b, bad_rows, ...]`` over ranks, in place, on
    the current stream (NCCL) -- the one collective of the path.  (Entries 5 and 6, the
    per-launch means, are meaningless after the sum: use :func:`means`.)"""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def means(stats):
    """-> (m_score, b_score) as the reference logs them (``utils.py:226, 324``)."""
    return stats[0] / stats[2], stats[1] / stats[3]


def get_cst_reward_sharded(model_res, data_gts, bcmr_scorer, bcmrscores=None, seq_per_img=20,
                           scb_captions=20, scb_baseline=1, use_eos=0, use_mixer=0, group=None):
    """``get_cst_reward`` on this rank's slice of a replicated global batch.

    -> (local rewards (n_local, T), global m_score, global b_score, (lo, hi) videos).
    A rank whose slice is empty (more ranks than videos in a short last batch) contributes zero
    statistics and still takes part in the all-reduce.
    Callers that already hold only their own slice call ``get_cst_reward(...,
    return_stats=True)`` and ``all_reduce_stats`` directly (that is what bench.py does).
    """
    from .reward import get_cst_reward
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    part = shard_batch(model_res, data_gts, seq_per_img, rank, world, bcmrscores=bcmrscores)
    rew, _, _, stats = get_cst_reward(part['model_res'], part['data_gts'], bcmr_scorer,
                                      bcmrscores=part.get('bcmrscores'), expand_feat=1,
                                      seq_per_img=seq_per_img, scb_captions=scb_captions,
                                      scb_baseline=scb_baseline, use_eos=use_eos, use_mixer=use_mixer,
                                      return_stats=True)
    all_reduce_stats(stats, group)
    m, b = means(stats)
    return rew, m, b, part['videos']
