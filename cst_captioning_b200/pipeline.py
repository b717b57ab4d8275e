"""Reward for batches that live in HOST memory: chunked upload overlapped with scoring.

The reference's data loader produces the reference rows on the host
(``dataloader.py:119-139``) and, in its own flow, even the sampled rows take a host
round trip (``train.py:185``).  When a caller holds the rows in pinned host memory,
uploading the whole batch and then scoring it serialises PCIe and the kernel; here the
batch is cut into slices of videos and slice i + 1 is uploaded on a copy stream while
slice i is scored on the compute stream, the rewards of slice i going back to the
caller's pinned output as soon as they exist.  Videos are independent, so slicing
changes no reward; the statistics sums simply add.
"""
from __future__ import annotations

import torch

from . import _lib
from .reward import CookedRefs, RefSet, _scorer_device, get_cst_reward


class HostRewardPipeline:
    """``get_cst_reward`` over pinned host tensors with copy/compute overlap.

    ``sampled``: pinned int tensor ``(videos * seq_per_img, T)``; ``refs``: pinned int
    tensor ``(videos, R, W)`` -- or a :class:`CookedRefs` batch (references staged once on
    the device: only the candidates are uploaded); optional ``bcmrscores`` pinned float64
    ``(videos, seq_per_img)``.  Keyword arguments are those of ``get_cst_reward``
    (``utils.py:229-238``) with ``expand_feat=1`` implied.
    """

    def __init__(self, scorer, seq_per_img=20, scb_captions=20, scb_baseline=1, use_eos=0, use_mixer=1,
                 slices=16):
        self.scorer = scorer
        self.device = _scorer_device(scorer)             # CiderD, or the Bleu / Rouge scorers
        self.kw = dict(expand_feat=1, seq_per_img=seq_per_img, scb_captions=scb_captions,
                       scb_baseline=scb_baseline, use_eos=use_eos, use_mixer=use_mixer, return_stats=True)
        self.seq_per_img = seq_per_img
        self.slices = max(1, int(slices))
        self.copy_stream = torch.cuda.Stream(self.device)
        self.out_stream = torch.cuda.Stream(self.device)
        self._bufs = None

    def _buffers(self, sampled, refs, bcmr, per):
        staged = isinstance(refs, CookedRefs)
        key = (per, sampled.shape[1], sampled.dtype, None if staged else tuple(refs.shape[1:]),
               None if staged else refs.dtype, bcmr is not None)
        if self._bufs is None or self._bufs[0] != key:
            def mk():
                return {'cand': torch.empty((per * self.seq_per_img, sampled.shape[1]), dtype=sampled.dtype,
                                            device=self.device),
                        'refs': None if staged else torch.empty((per,) + tuple(refs.shape[1:]), dtype=refs.dtype,
                                                                device=self.device),
                        'bcmr': torch.empty((per, self.seq_per_img), dtype=torch.float64, device=self.device)
                        if bcmr is not None else None,
                        'uploaded': torch.cuda.Event(), 'consumed': torch.cuda.Event(),
                        'scored': torch.cuda.Event(), 'used': False}
            self._bufs = (key, [mk(), mk()], torch.zeros(_lib.N_STATS, dtype=torch.float64, device=self.device))
        return self._bufs[1]

    def __call__(self, sampled, refs, bcmrscores=None, out=None):
        """-> (rewards float32 pinned host tensor (N,), stats float64 CUDA tensor (5,)).
        The caller synchronises (``torch.cuda.synchronize()`` or the returned tensors'
        streams) before reading ``out``."""
        s = self.seq_per_img
        staged = isinstance(refs, CookedRefs)
        n_videos = refs.n_videos if staged else refs.shape[0]
        if staged:
            index = refs.index if refs.index is not None else torch.arange(n_videos, device=self.device)
        if sampled.shape[0] != n_videos * s:
            raise ValueError('sampled must hold seq_per_img rows per video')
        if out is None:
            out = torch.empty(sampled.shape[0], dtype=torch.float32).pin_memory()
        if n_videos == 0:                                # an empty shard: nothing to score, zero statistics
            return out, torch.zeros(_lib.N_STATS, dtype=torch.float64, device=self.device)
        per = (n_videos + self.slices - 1) // self.slices
        bufs = self._buffers(sampled, refs, bcmrscores, per)
        main = torch.cuda.current_stream(self.device)
        self.copy_stream.wait_stream(main)
        total = self._bufs[2]
        total.zero_()
        done = []
        for i, lo in enumerate(range(0, n_videos, per)):
            hi = min(n_videos, lo + per)
            b = bufs[i & 1]
            with torch.cuda.stream(self.copy_stream):
                if b['used']:
                    self.copy_stream.wait_event(b['consumed'])      # the slice two back has been scored
                cand = b['cand'][:(hi - lo) * s]
                cand.copy_(sampled[lo * s:hi * s], non_blocking=True)
                if staged:
                    rr = refs.batch(index[lo:hi])
                else:
                    rr = b['refs'][:hi - lo]
                    rr.copy_(refs[lo:hi], non_blocking=True)
                bc = None
                if bcmrscores is not None:
                    bc = b['bcmr'][:hi - lo]
                    bc.copy_(bcmrscores[lo:hi], non_blocking=True)
                b['uploaded'].record(self.copy_stream)
            main.wait_event(b['uploaded'])
            rew, _, _, stats = get_cst_reward(cand, rr, self.scorer, bcmrscores=bc, **self.kw)
            total += stats
            b['consumed'].record(main)
            b['used'] = True
            b['scored'].record(main)
            with torch.cuda.stream(self.out_stream):
                self.out_stream.wait_event(b['scored'])
                out[lo * s:hi * s].copy_(rew[:, 0], non_blocking=True)
                rew.record_stream(self.out_stream)
            done.append(rew)
        total[5] = total[0] / total[2]              # means of the whole batch (the slices' own means were summed)
        total[6] = total[1] / total[3]
        main.wait_stream(self.out_stream)
        for b in bufs:
            b['used'] = False
        main.wait_stream(self.copy_stream)
        return out, total.clone()
