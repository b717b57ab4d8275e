"""``RewardCriterion`` fused on the device.

Mirror of ``model.RewardCriterion`` (``model.py:7-23``; called at
``train.py:206-210``): ``loss = sum(-logprobs * reward * mask) / sum(mask)`` with the
mask of ``model.py:15-18`` (ones shifted right by one so the ``<end>`` token is
counted).  One streaming kernel each for the forward and the backward; the reward
column produced by :func:`get_cst_reward` (an ``expand``-ed ``(N, T)`` view) is read
through its strides, so the broadcast is never materialised.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import lib, check

_scratch = {}


def _scratch_for(device):
    key = (device, torch.cuda.current_stream(device).cuda_stream)     # never shared by overlapping launches
    buf = _scratch.get(key)
    if buf is None:
        buf = torch.zeros(int(lib.cst_criterion_scratch_bytes()), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def _seq_dtype(seq):
    if seq.dtype == torch.int64:
        return _lib.DTYPE_I64
    if seq.dtype == torch.int32:
        return _lib.DTYPE_I32
    raise ValueError('seq must be int32 or int64, got %s' % seq.dtype)


def _reward_strides(reward, n, t):
    if reward.dim() == 1:
        if reward.shape[0] != n:
            raise ValueError('reward must hold one value per caption')
        return reward.stride(0), 0
    if tuple(reward.shape) != (n, t):
        raise ValueError('reward must be (N,) or (N, T)')
    return reward.stride(0), reward.stride(1)


class _RewardLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, seq, logprobs, reward):
        if not (seq.is_cuda and logprobs.is_cuda and reward.is_cuda):
            raise ValueError('RewardCriterion runs on CUDA tensors only (no CPU fallback)')
        n, t = seq.shape
        if tuple(logprobs.shape) != (n, t):
            raise ValueError('logprobs must have the shape of seq')
        if seq.stride(1) != 1:
            seq = seq.contiguous()
        lp = logprobs.detach()
        if lp.dtype != torch.float32 or lp.stride(1) != 1:
            lp = lp.float().contiguous()
        rw = reward.detach()
        if rw.dtype != torch.float32:
            rw = rw.float()
        rs, cs = _reward_strides(rw, n, t)
        dev = seq.device
        out = torch.empty(2, dtype=torch.float32, device=dev)          # [loss, sum(mask)]
        scratch = _scratch_for(dev)
        check(lib.cst_reward_criterion_fwd(
            seq.data_ptr(), _seq_dtype(seq), seq.stride(0), lp.data_ptr(), lp.stride(0),
            rw.data_ptr(), rs, cs, n, t, out.data_ptr(), out.data_ptr() + 4,
            scratch.data_ptr(), scratch.numel(), dev.index, torch.cuda.current_stream(dev).cuda_stream))
        ctx.save_for_backward(seq, rw, out)
        ctx.strides = (rs, cs)
        ctx.lp_dtype = logprobs.dtype
        return out[0]

    @staticmethod
    def backward(ctx, grad_out):
        seq, rw, out = ctx.saved_tensors
        n, t = seq.shape
        dev = seq.device
        rs, cs = ctx.strides
        go = grad_out.detach().to(torch.float32).reshape(1).contiguous()
        grad = torch.empty((n, t), dtype=torch.float32, device=dev)
        check(lib.cst_reward_criterion_bwd(
            seq.data_ptr(), _seq_dtype(seq), seq.stride(0), rw.data_ptr(), rs, cs, out.data_ptr() + 4,
            go.data_ptr(), n, t, grad.data_ptr(), grad.stride(0), dev.index,
            torch.cuda.current_stream(dev).cuda_stream))
        return None, grad.to(ctx.lp_dtype), None


class RewardCriterion(torch.nn.Module):
    """Drop-in for ``model.RewardCriterion``: ``forward(seq, logprobs, reward)``.
    ``reward`` may be the ``(N, T)`` tensor of the reference or the ``(N,)`` column."""

    def forward(self, seq, logprobs, reward):
        return _RewardLoss.apply(seq, logprobs, reward)
