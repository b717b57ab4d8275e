"""GT consensus scores on the device: the CIDEr column of ``compute_scores.py``.

``compute_scores.py:75-103`` (``make compute_evalscores``, run with
``--remove_in_ref``, ``Makefile:118-120``) scores caption i of every video against the
video's other captions with coco-caption's ``Cider`` in *corpus* mode: the document
frequency is recomputed from the references of that very call and
``ref_len = log(number of videos)``.  The result, ``(videos, seq_per_img)`` float64, is
what the data loader later hands to the reward path as ``bcmrscores``
(``dataloader.py:66-73, 142-144``) for the SCB baseline and the WXE weights.

Here each of the ``seq_per_img`` passes builds its corpus-mode df table on the device
(``DfTable.from_rows``) and scores column i against the rest with the same fused kernel
as the RL path.

Difference in input, not in arithmetic: the reference scores the raw caption *words*
of the cocofmt json, this module scores token-id rows.  They agree when the rows hold
every word of the caption (a vocabulary without ``<unk>`` folding, captions of at most
32 tokens); rows that were truncated or UNK-folded score the folded caption.
"""
from __future__ import annotations

import torch

from .ciderd import CiderD
from .df_table import DfTable
from .reward import RefSet


def consensus_scores(refs, seq_per_img=None, remove_in_ref=True, use_eos=0):
    """``refs``: CUDA int tensor ``(videos, R, W)`` of reference rows (``[<start>, w..,
    <end>, pad]`` label rows or bare word rows).  -> float64 CUDA tensor
    ``(videos, seq_per_img)``: column i = CIDEr-D of caption i against the others."""
    if not (torch.is_tensor(refs) and refs.is_cuda and refs.dim() == 3):
        raise ValueError('refs must be a CUDA tensor (videos, refs, width)')
    n_videos, n_refs, width = refs.shape
    seq_per_img = n_refs if seq_per_img is None else int(seq_per_img)
    if not 1 <= seq_per_img <= n_refs:
        raise ValueError('seq_per_img must lie in [1, refs per video]')
    if remove_in_ref and n_refs < 2:
        raise ValueError('remove_in_ref needs at least two captions per video')
    out = torch.empty(n_videos, seq_per_img, dtype=torch.float64, device=refs.device)
    shared = None
    if not remove_in_ref:                       # one corpus, one table for every pass
        shared = (CiderD(df=DfTable.from_rows(refs, use_eos=use_eos)),
                  RefSet(refs.reshape(n_videos * n_refs, width), per_video=n_refs))
    for i in range(seq_per_img):
        if remove_in_ref:                       # compute_scores.py:86-87
            others = torch.cat([refs[:, :i], refs[:, i + 1:]], dim=1).contiguous()
            scorer = CiderD(df=DfTable.from_rows(others, use_eos=use_eos))
            refset = RefSet(others.reshape(n_videos * (n_refs - 1), width), per_video=n_refs - 1)
        else:
            scorer, refset = shared
        hyp = refs[:, i].contiguous()
        out[:, i] = scorer.score_rows(hyp, refset, 1, use_eos=use_eos)
        if remove_in_ref:
            torch.cuda.current_stream(refs.device).synchronize()   # table is freed with `scorer`
            scorer.table.close()
    return out
