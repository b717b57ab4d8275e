"""GT consensus scores on the device: the CIDEr, Bleu_4 and ROUGE_L columns of ``compute_scores.py``.

``compute_scores.py:75-103`` (``make compute_evalscores``, run with
``--remove_in_ref``, ``Makefile:118-120``) scores caption i of every video against the
video's other captions with coco-caption's ``Cider`` in *corpus* mode: the document
frequency is recomputed from the references of that very call and
``ref_len = log(number of videos)``.  The result, ``(videos, seq_per_img)`` float64, is
what the data loader later hands to the reward path as ``bcmrscores``
(``dataloader.py:66-73, 142-144``) for the SCB baseline and the WXE weights.

The twenty corpus-mode tables differ from the df of ALL captions only where caption i is a
video's sole holder of an n-gram: ``df_i(g) = df_all(g) - #{videos whose caption i alone holds
g}``.  So the table is built ONCE from all rows, one more pass over the rows counts the sole holders
per (caption, n-gram) (``cst_df_sole_holders``), and pass i switches the table's idf to
``log V - log max(1, df_all - excl_i)`` with one streaming kernel (``cst_df_leave_one_out``) before
scoring column i against the rest with the same fused kernel as the RL path.  Videos whose captions
exceed one shared-memory table (more than 32 captions or 1,100 n-gram instances) fall back to one
device-built table per pass.

The ``Bleu_4`` and ``ROUGE_L`` columns (``compute_scores.py:67-68``; ``METEOR`` is a Java program) need
no corpus statistics: pass i is one launch of ``k_bleu`` / ``k_rouge`` on caption i against the others.
``evalscores`` computes the columns wanted and ``write_evalscores_pkl`` stores them the way
``compute_scores.py:105-108`` does, for ``dataloader.py:66-73`` to read (``--eval_metric`` picks the key).

Difference in input, not in arithmetic: the reference scores the raw caption *words*
of the cocofmt json, this module scores token-id rows.  They agree when the rows hold
every word of the caption (a vocabulary without ``<unk>`` folding, captions of at most
32 tokens); rows that were truncated or UNK-folded score the folded caption.
"""
from __future__ import annotations

import torch

from ._lib import lib, check
from .ciderd import CiderD
from .df_table import DfTable, _row_dtype
from .reward import RefSet


def consensus_scores(refs, seq_per_img=None, remove_in_ref=True, use_eos=0, metric='CIDEr'):
    """``refs``: CUDA int tensor ``(videos, R, W)`` of reference rows (``[<start>, w..,
    <end>, pad]`` label rows or bare word rows).  -> float64 CUDA tensor
    ``(videos, seq_per_img)``: column i = score of caption i against the others under ``metric``
    ('CIDEr', 'Bleu_4' or 'ROUGE_L': the keys of ``compute_scores.py:67-68``)."""
    if metric not in ('CIDEr', 'Bleu_4', 'ROUGE_L'):
        raise ValueError("metric must be 'CIDEr', 'Bleu_4' or 'ROUGE_L' (METEOR is a Java program)")
    if not (torch.is_tensor(refs) and refs.is_cuda and refs.dim() == 3):
        raise ValueError('refs must be a CUDA tensor (videos, refs, width)')
    n_videos, n_refs, width = refs.shape
    seq_per_img = n_refs if seq_per_img is None else int(seq_per_img)
    if not 1 <= seq_per_img <= n_refs:
        raise ValueError('seq_per_img must lie in [1, refs per video]')
    if remove_in_ref and n_refs < 2:
        raise ValueError('remove_in_ref needs at least two captions per video')
    refs = refs.contiguous()
    dev = refs.device
    out = torch.empty(n_videos, seq_per_img, dtype=torch.float64, device=dev)
    if n_videos == 0:
        return out
    if metric != 'CIDEr':
        from .metrics import Bleu, Rouge
        scorer = Bleu(device=dev) if metric == 'Bleu_4' else Rouge(device=dev)
        for i in range(seq_per_img):                          # compute_scores.py:81-92, no corpus statistics
            if remove_in_ref:
                others = torch.cat([refs[:, :i], refs[:, i + 1:]], dim=1).contiguous()
            else:
                others = refs
            refset = RefSet(others.reshape(-1, width), per_video=others.shape[1])
            out[:, i] = scorer.score_rows(refs[:, i].contiguous(), refset, 1, use_eos=use_eos)
        return out
    table = DfTable.from_rows(refs, use_eos=use_eos)          # df of ALL captions, ref_len = log(videos)
    scorer = CiderD(df=table)
    flat = refs.reshape(n_videos * n_refs, width)
    if not remove_in_ref:                                     # one corpus for every pass
        refset = RefSet(flat, per_video=n_refs)
        for i in range(seq_per_img):
            out[:, i] = scorer.score_rows(refs[:, i], refset, 1, use_eos=use_eos)
        torch.cuda.current_stream(dev).synchronize()
        table.close()
        return out
    stream = torch.cuda.current_stream(dev).cuda_stream
    excl = None
    if n_refs <= 32:
        excl = torch.zeros((n_refs, table.capacity), dtype=torch.int32, device=dev)
        too_large = torch.zeros(1, dtype=torch.int32, device=dev)
        check(lib.cst_df_sole_holders(table.handle, flat.data_ptr(), _row_dtype(flat), flat.shape[0], width,
                                      flat.stride(0), n_refs, n_videos, int(use_eos), excl.data_ptr(),
                                      too_large.data_ptr(), stream))
        if int(too_large.item()) != 0:
            excl = None
    for i in range(seq_per_img):                              # compute_scores.py:81-92
        others = torch.cat([refs[:, :i], refs[:, i + 1:]], dim=1).contiguous()
        refset = RefSet(others.reshape(n_videos * (n_refs - 1), width), per_video=n_refs - 1)
        if excl is not None:
            check(lib.cst_df_leave_one_out(table.handle, excl.data_ptr(), i, stream))
            out[:, i] = scorer.score_rows(refs[:, i], refset, 1, use_eos=use_eos)
        else:                                                 # a table of its own for this pass
            t_i = DfTable.from_rows(others, use_eos=use_eos)
            out[:, i] = CiderD(df=t_i).score_rows(refs[:, i].contiguous(), refset, 1, use_eos=use_eos)
            torch.cuda.current_stream(dev).synchronize()
            t_i.close()
    torch.cuda.current_stream(dev).synchronize()              # the table is freed with `scorer`
    table.close()
    return out


def evalscores(refs, seq_per_img=None, remove_in_ref=True, use_eos=0, metrics=('Bleu_4', 'ROUGE_L', 'CIDEr')):
    """The table ``compute_scores.py:75-103`` builds: {metric: float64 CUDA tensor (videos, seq_per_img)}."""
    return {m: consensus_scores(refs, seq_per_img, remove_in_ref, use_eos, metric=m) for m in metrics}


def write_evalscores_pkl(path, scores):
    """``compute_scores.py:105-108``: the dict of (videos, seq_per_img) float64 arrays, pickled with the
    highest protocol -- what ``dataloader.py:66-73`` loads as ``bcmrscores_pkl``."""
    import pickle
    host = {m: (v.cpu().numpy() if torch.is_tensor(v) else v) for m, v in scores.items()}
    with open(path, 'wb') as fh:
        pickle.dump(host, fh, protocol=pickle.HIGHEST_PROTOCOL)
    return host
