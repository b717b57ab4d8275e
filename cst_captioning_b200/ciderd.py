"""``CiderD`` with the signature of ``pyciderevalcap.ciderD.ciderD.CiderD``.

Construction site ``train.py:121`` (``CiderD(df=opt.train_cached_tokens)``),
call sites ``utils.py:72, 208, 274`` (``compute_score(gts, res)``).  The object
owns a device-resident :class:`DfTable`; the reward functions in
``cst_captioning_b200.reward`` use it directly on token rows, while
``compute_score`` keeps the string protocol so the reference's unmodified
``utils.py`` can drive the GPU path.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from ._lib import lib, check
from .df_table import DfTable, _row_dtype


def _parse(sentence, width):
    toks = sentence.split()
    if len(toks) > width:
        raise ValueError('caption of %d tokens exceeds the kernel row width %d' % (len(toks), width))
    try:
        ids = [int(t) for t in toks]
    except ValueError:
        raise ValueError('the GPU scorer works on token-id strings (utils.array_to_str output); '
                         'got %r' % sentence)
    if 1 in ids or 0 in ids[:-1]:
        raise ValueError('not an array_to_str string (<start> or an inner <end>): %r' % sentence)
    return ids


class CiderD:
    """CIDEr-D scorer in cached-df mode (n = 4, sigma = 6.0, the defaults the
    reference constructs it with)."""

    def __init__(self, n=4, sigma=6.0, df=None, device=None):
        if n != 4 or float(sigma) != 6.0:
            raise ValueError('the sm_100a kernels implement n=4, sigma=6.0 (the reference defaults)')
        if isinstance(df, DfTable):
            self.table = df
        elif isinstance(df, (str, bytes)):
            self.table = DfTable.from_pickle(df, device)
        elif isinstance(df, tuple) and len(df) == 2:
            self.table = DfTable.from_dict(df[0], df[1], device)
        else:
            raise ValueError("df must be a *_ciderdf.pkl path, a (mapping, ref_len) pair or a DfTable; "
                             "the 'corpus' mode lives in cst_captioning_b200.consensus")
        self._n = n
        self._sigma = sigma

    # -- tensor level ----------------------------------------------------------
    def score_rows(self, cand, refs, cands_per_video, use_eos=0, greedy=None):
        """CIDEr-D of token rows.  ``cand`` (N, T) CUDA int tensor, ``refs`` a
        :class:`~cst_captioning_b200.reward.RefSet`; candidate i is scored against
        video ``i // cands_per_video``.  -> float64 CUDA tensor (N,), or (2N,) with
        ``greedy`` rows appended (``utils.py:180-183`` order)."""
        dev = self.table.device
        n = cand.shape[0]
        out = torch.empty(n * (2 if greedy is not None else 1), dtype=torch.float64, device=dev)
        check(lib.cst_cider_scores(
            self.table.handle, cand.data_ptr(), greedy.data_ptr() if greedy is not None else None,
            _row_dtype(cand), n, cand.shape[1], cand.stride(0),
            refs.rows.data_ptr(), _row_dtype(refs.rows), refs.rows.shape[1], refs.rows.stride(0),
            refs.offsets.data_ptr() if refs.offsets is not None else None, refs.max_refs,
            n // cands_per_video, cands_per_video, int(use_eos), out.data_ptr(),
            torch.cuda.current_stream(dev).cuda_stream))
        return out

    # -- string level (the reference protocol) -----------------------------------
    def compute_score(self, gts, res):
        """``gts``: {image_id: [ref strings]}, ``res``: [{'image_id', 'caption': [str]}]
        -> (mean, float64 ndarray).  Strings are space-joined token ids as
        ``utils.array_to_str`` emits them (``utils.py:135-152``)."""
        from .reward import RefSet
        width = _lib.MAX_ROW_WIDTH
        n = len(res)
        if n == 0:
            return float('nan'), np.zeros(0)
        cand = np.full((n, width), -1, dtype=np.int32)
        groups, ref_rows, counts = [], [], []
        last = None
        for i, entry in enumerate(res):
            hyp = entry['caption']
            refs = gts[entry['image_id']]
            assert type(hyp) is list and len(hyp) == 1
            assert type(refs) is list and len(refs) > 0
            ids = _parse(hyp[0], width)
            cand[i, :len(ids)] = ids
            if refs is last:
                groups[-1] += 1
                continue
            last = refs
            groups.append(1)
            counts.append(len(refs))
            for r in refs:
                row = np.full(width, -1, dtype=np.int32)
                ids = _parse(r, width)
                row[:len(ids)] = ids
                ref_rows.append(row)
        # The strings are already canonical: rows are the ids followed by the padding
        # id -1 (ends a caption without being emitted) and are scored with use_eos=1,
        # which keeps a trailing <end> = 0 exactly when the string has one.
        dev = self.table.device
        per = groups[0] if len(set(groups)) == 1 else None
        if per is None:        # uneven sharing of ref lists: one "video" per hypothesis
            expanded, counts2 = [], []
            at = 0
            for g, c in zip(groups, counts):
                for _ in range(g):
                    expanded.extend(ref_rows[at:at + c])
                    counts2.append(c)
                at += c
            ref_rows, counts, per = expanded, counts2, 1
        rows = torch.from_numpy(np.stack(ref_rows)).to(dev)
        if len(set(counts)) == 1:
            refset = RefSet(rows, per_video=counts[0])
        else:
            refset = RefSet(rows, offsets=np.concatenate([[0], np.cumsum(counts)]))
        scores = self.score_rows(torch.from_numpy(cand).to(dev), refset, per, use_eos=1)
        arr = scores.cpu().numpy()
        return np.mean(arr), arr


class Cider:
    """``pycocoevalcap.cider.cider.Cider`` (corpus mode) with the string protocol of
    ``compute_scores.py:67-68, 92``: ``compute_score(gts, res)`` with both dicts keyed by
    image id and ``res[id]`` a one-element list.  The document frequency is rebuilt on the
    device from the references of THIS call (``ref_len = log(number of images)``), then
    every hypothesis is scored with the same kernel as the RL path.

    The strings must be space-joined token ids (what ``utils.array_to_str`` emits); the
    reference's ``compute_scores.py`` feeds raw words, which have to be mapped to ids
    first (any injective word -> id map gives the same scores).
    """

    def __init__(self, n=4, sigma=6.0, device=None):
        if n != 4 or float(sigma) != 6.0:
            raise ValueError('the sm_100a kernels implement n=4, sigma=6.0 (the reference defaults)')
        self._device = torch.device(device if device is not None else 'cuda')

    def compute_score(self, gts, res):
        from .reward import RefSet
        assert gts.keys() == res.keys()
        ids = list(gts.keys())
        width = _lib.MAX_ROW_WIDTH
        counts, ref_rows = [], []
        cand = np.full((len(ids), width), -1, dtype=np.int32)
        for i, key in enumerate(ids):
            hyp, refs = res[key], gts[key]
            assert type(hyp) is list and len(hyp) == 1
            assert type(refs) is list and len(refs) > 0
            toks = _parse(hyp[0], width)
            cand[i, :len(toks)] = toks
            counts.append(len(refs))
            for r in refs:
                row = np.full(width, -1, dtype=np.int32)
                toks = _parse(r, width)
                row[:len(toks)] = toks
                ref_rows.append(row)
        dev = self._device
        rows = torch.from_numpy(np.stack(ref_rows)).to(dev)
        offsets = np.concatenate([[0], np.cumsum(counts)])
        table = DfTable.from_rows(rows, video_offsets=offsets, use_eos=1)
        scorer = CiderD(df=table)
        refset = RefSet(rows, per_video=counts[0]) if len(set(counts)) == 1 else RefSet(rows, offsets=offsets)
        arr = scorer.score_rows(torch.from_numpy(cand).to(dev), refset, 1, use_eos=1).cpu().numpy()
        table.close()
        return np.mean(arr), arr
