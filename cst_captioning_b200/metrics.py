"""``Bleu`` and ``Rouge`` with the signatures of ``pycocoevalcap.bleu.bleu.Bleu`` /
``pycocoevalcap.rouge.rouge.Rouge``: the reward scorers ``--eval_metric Bleu_4`` / ``ROUGE_L`` select
(``train.py:119-124``), on the GPU.

``compute_score(gts, res)`` keeps the dict protocol of the call sites ``utils.py:72, 208, 274`` (ids ->
lists of strings), so the reference's unmodified ``utils.py`` can drive these classes;
``score_rows`` works on token rows and is what ``cst_captioning_b200.reward`` calls:
``get_cst_reward`` / ``get_self_critical_reward`` accept these scorers wherever they accept ``CiderD``
(for ``Bleu`` the BLEU-4 sentence scores count, ``utils.py:210-213, 276-279``).  ``Meteor`` is a Java
program in the reference's checkout and has no counterpart.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib
from ._lib import lib, check
from .ciderd import WordIds, _row
from .df_table import _row_dtype


def _device(device):
    dev = torch.device(device if device is not None else 'cuda')
    if dev.index is None:
        dev = torch.device('cuda', torch.cuda.current_device())
    return dev


def _row_args(cand, greedy, refs, cands_per_video, use_eos, dev):
    n = cand.shape[0]
    if cand.device != dev or refs.rows.device != dev or (greedy is not None and greedy.device != dev):
        raise ValueError('rows must live on the scorer\'s device %s' % dev)
    if greedy is not None and (greedy.shape != cand.shape or greedy.dtype != cand.dtype or
                               greedy.stride(0) != cand.stride(0)):
        raise ValueError('greedy rows must share the geometry of the sampled rows')
    if n % cands_per_video or n // cands_per_video > refs.n_videos:
        raise ValueError('caption count %d does not match %d videos x %d' % (n, refs.n_videos, cands_per_video))
    return [cand.data_ptr(), greedy.data_ptr() if greedy is not None else None, _row_dtype(cand), n,
            cand.shape[1], cand.stride(0), refs.rows.data_ptr(), _row_dtype(refs.rows), refs.rows.shape[1],
            refs.rows.stride(0), refs.offsets.data_ptr() if refs.offsets is not None else None, refs.max_refs,
            n // cands_per_video, cands_per_video, int(use_eos)]


class _RowScorer:
    """What the reward functions need of a scorer that is not ``CiderD``."""

    def __init__(self, device=None, on_long='error'):
        self.device = _device(device)
        self._on_long = on_long          # captions beyond the 32 tokens of a kernel row: 'error' or 'truncate'
        self.truncated = 0

    def _strings_to_rows(self, gts, res, sep):
        """dict protocol -> (candidate rows, RefSet): every distinct word of the call is interned to an
        id; each entry is a "video" of its own (one candidate against its list of references)."""
        from .reward import RefSet
        assert gts.keys() == res.keys()
        keys = list(gts.keys())
        width = _lib.MAX_ROW_WIDTH
        words = WordIds(max_tokens=width, on_long=self._on_long)
        cand = np.full((len(keys), width), -1, dtype=np.int32)
        counts, ref_rows = [], []
        for i, key in enumerate(keys):
            hyp, refs = res[key], gts[key]
            assert type(hyp) is list and len(hyp) == 1
            assert type(refs) is list and len(refs) > 0
            toks = words.ids(hyp[0], sep=sep)
            cand[i, :len(toks)] = toks
            counts.append(len(refs))
            for r in refs:
                ref_rows.append(_row(words.ids(r, sep=sep), width))
        self.truncated += words.truncated
        dev = self.device
        rows = torch.from_numpy(np.stack(ref_rows)).to(dev)
        if len(set(counts)) == 1:
            refset = RefSet(rows, per_video=counts[0])
        else:
            refset = RefSet(rows, offsets=np.concatenate([[0], np.cumsum(counts)]))
        return torch.from_numpy(cand).to(dev), refset


class Bleu(_RowScorer):
    """Sentence and corpus BLEU-1..4, ``BleuScorer.compute_score(option='closest')``."""

    def __init__(self, n=4, device=None, on_long='error'):
        if n != 4:
            raise ValueError('the sm_100a kernel implements n=4 (the reference default)')
        super().__init__(device, on_long)
        self._n = n

    def bleu_rows(self, cand, refs, cands_per_video, use_eos=0, greedy=None, comps=False, bad=None):
        """-> float64 CUDA tensor (4, n_total) of sentence BLEU-1..4 (and, with ``comps``, the int32
        (n_total, 6) statistics testlen, reflen, correct[4]).  ``bad``: int32 CUDA tensor (1,) that the
        number of rows holding ids outside [-1, 65534] is added to."""
        dev = self.device
        n_total = cand.shape[0] * (2 if greedy is not None else 1)
        out = torch.empty((4, n_total), dtype=torch.float64, device=dev)
        stats = torch.empty((n_total, 6), dtype=torch.int32, device=dev) if comps else None
        check(lib.cst_bleu_scores(*_row_args(cand, greedy, refs, cands_per_video, use_eos, dev), out.data_ptr(),
                                  stats.data_ptr() if comps else None, bad.data_ptr() if bad is not None else None, dev.index,
                                  torch.cuda.current_stream(dev).cuda_stream))
        return (out, stats) if comps else out

    def score_rows(self, cand, refs, cands_per_video, use_eos=0, greedy=None, bad=None):
        """BLEU-4 of token rows -> float64 CUDA tensor (n_total,): what the reward reads."""
        return self.bleu_rows(cand, refs, cands_per_video, use_eos, greedy, bad=bad)[3]

    def compute_score(self, gts, res):
        """-> ([BLEU-1..4 of the corpus], [4 lists of sentence scores])."""
        if len(res) == 0:
            return [0.0] * 4, [[] for _ in range(4)]
        cand, refset = self._strings_to_rows(gts, res, None)
        sent, comps = self.bleu_rows(cand, refset, 1, use_eos=0, comps=True)
        c = comps.to(torch.int64)
        testlen, reflen = int(c[:, 0].sum()), int(c[:, 1].sum())
        small, tiny = 1e-9, 1e-15
        bleus, bleu = [], 1.
        for k in range(4):
            correct = int(c[:, 2 + k].sum())
            guess = int((c[:, 0] - k).clamp_(min=0).sum())
            bleu *= float(correct + tiny) / (guess + small)
            bleus.append(bleu ** (1. / (k + 1)))
        ratio = (testlen + tiny) / (reflen + small)
        if ratio < 1:
            bleus = [b * math.exp(1 - 1 / ratio) for b in bleus]
        return bleus, [row.tolist() for row in sent.cpu()]


class Rouge(_RowScorer):
    """ROUGE-L F-measure (beta = 1.2) of ``Rouge.calc_score``."""

    def __init__(self, device=None, on_long='error'):
        super().__init__(device, on_long)
        self.beta = 1.2

    def score_rows(self, cand, refs, cands_per_video, use_eos=0, greedy=None, bad=None):
        """ROUGE-L of token rows -> float64 CUDA tensor (n_total,); ``bad`` as in ``Bleu.bleu_rows``."""
        dev = self.device
        out = torch.empty(cand.shape[0] * (2 if greedy is not None else 1), dtype=torch.float64, device=dev)
        check(lib.cst_rouge_scores(*_row_args(cand, greedy, refs, cands_per_video, use_eos, dev), out.data_ptr(),
                                   bad.data_ptr() if bad is not None else None,
                                   dev.index, torch.cuda.current_stream(dev).cuda_stream))
        return out

    def compute_score(self, gts, res):
        """-> (mean, float64 ndarray).  Sentences are cut with ``split(" ")`` like the scorer's."""
        if len(res) == 0:
            return float('nan'), np.zeros(0)
        cand, refset = self._strings_to_rows(gts, res, ' ')
        arr = self.score_rows(cand, refset, 1, use_eos=0).cpu().numpy()
        return np.mean(arr), arr
