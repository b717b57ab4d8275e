"""Reward functions of the RL loop on the GPU, with the reference's signatures.

Mirrors ``utils.get_cst_reward`` (``utils.py:229-324``) and
``utils.get_self_critical_reward`` (``utils.py:169-226``): same argument names,
meaning and error behaviour; the work happens in one launch of the sm_100a
kernels behind the C ABI (``include/cst_reward.h``) and the results stay on the
device:

* ``rewards``  -> float32 CUDA tensor ``(N, T)``, an ``expand``-ed view of the
  per-caption column (the reference's ``np.repeat(...)``, ``utils.py:224, 322``,
  followed by ``torch.from_numpy(reward).float().cuda()``, ``train.py:209``);
* ``m_score``, ``b_score``/``g_score`` -> 0-dim float64 CUDA tensors
  (``float(x)`` at log time, ``train.py:229-233``).

``model_res`` / ``greedy_res`` may be CUDA tensors (no host round trip, the
``.cpu().numpy()`` of ``train.py:185,198-199`` is simply dropped) or numpy arrays
(drop-in for the unmodified call sites; they are uploaded).  ``data_gts`` may be
the reference's list of per-video ``(ncap, W)`` arrays (``dataloader.py:138-139``),
a CUDA tensor ``(B, R, W)``, or a :class:`RefSet` kept on the device.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from ._lib import lib, check
from .ciderd import CiderD
from .df_table import _row_dtype


class RefSet:
    """Reference rows resident on the device: ``rows`` (total, W) CUDA int tensor,
    video v owning rows ``offsets[v]:offsets[v+1]`` (None = uniform ``per_video``)."""

    def __init__(self, rows, offsets=None, per_video=None):
        if not (torch.is_tensor(rows) and rows.is_cuda and rows.dim() == 2):
            raise ValueError('RefSet rows must be a 2-D CUDA tensor')
        if rows.stride(1) != 1:
            rows = rows.contiguous()
        self.rows = rows
        if offsets is None:
            if not per_video or per_video < 1 or rows.shape[0] % per_video:
                raise ValueError('uniform RefSet needs per_video dividing the row count')
            self.offsets = None
            self.max_refs = int(per_video)
            self.n_videos = rows.shape[0] // per_video
        else:
            off = np.ascontiguousarray(offsets.cpu().numpy() if torch.is_tensor(offsets) else offsets,
                                       dtype=np.int64)
            if off[0] != 0 or off[-1] != rows.shape[0] or np.any(np.diff(off) < 0):
                raise ValueError('offsets must rise from 0 to the row count')
            self.offsets = torch.from_numpy(off).to(rows.device)
            self.max_refs = int(np.diff(off).max()) if len(off) > 1 else 0
            self.n_videos = len(off) - 1

    def __len__(self):
        return self.n_videos

    def __getitem__(self, videos):
        """A contiguous slice of the videos (what ``dist.shard_batch`` takes of ``data_gts``)."""
        if not isinstance(videos, slice):
            raise TypeError('RefSet supports slices of videos')
        lo, hi, step = videos.indices(self.n_videos)
        if step != 1:
            raise ValueError('RefSet slices must be contiguous')
        hi = max(hi, lo)
        if self.offsets is None:
            return RefSet(self.rows[lo * self.max_refs:hi * self.max_refs], per_video=self.max_refs)
        off = self.offsets[lo:hi + 1].cpu().numpy()
        return RefSet(self.rows[int(off[0]):int(off[-1])], offsets=off - off[0])

    def take(self, video_ids):
        """The rows of the named videos, in order, gathered on the device -> RefSet."""
        ids = np.ascontiguousarray(video_ids, dtype=np.int64)
        if self.offsets is None:
            r = self.max_refs
            rows = (ids[:, None] * r + np.arange(r)[None, :]).reshape(-1)
            return RefSet(self.rows.index_select(0, torch.from_numpy(rows).to(self.rows.device)), per_video=r)
        off = self.offsets.cpu().numpy()
        counts = off[ids + 1] - off[ids]
        new_off = np.concatenate([[0], np.cumsum(counts)])
        rows = np.repeat(off[ids] - new_off[:-1], counts) + np.arange(new_off[-1])
        return RefSet(self.rows.index_select(0, torch.from_numpy(rows).to(self.rows.device)), offsets=new_off)

    @classmethod
    def from_list(cls, data_gts, device):
        counts = [len(g) for g in data_gts]
        if not counts:                                   # an empty shard: no videos, no rows
            return cls(torch.zeros((0, 1), dtype=torch.int32, device=device), offsets=np.zeros(1, dtype=np.int64))
        rows = np.concatenate([np.asarray(g).reshape(len(g), -1) for g in data_gts], axis=0)
        if rows.dtype.kind not in 'iu':
            if np.any(rows != np.floor(rows)):
                raise ValueError('reference rows must hold integer token ids')
        rows = _narrow(np.ascontiguousarray(rows, dtype=np.int64))
        t = torch.from_numpy(rows).to(device)
        if len(set(counts)) == 1:
            return cls(t, per_video=counts[0])
        return cls(t, offsets=np.concatenate([[0], np.cumsum(counts)]))

    @classmethod
    def coerce(cls, data_gts, device):
        if isinstance(data_gts, (cls, CookedRefs)):
            return data_gts
        if torch.is_tensor(data_gts):
            if data_gts.dim() != 3:
                raise ValueError('a reference tensor must be (videos, refs, width)')
            b, r, w = data_gts.shape
            return cls(data_gts.to(device).reshape(b * r, w), per_video=r)
        return cls.from_list(data_gts, device)


class RefStore:
    """The whole label matrix kept on the device.

    The reference's loader re-reads the references of every batch video from HDF5 and
    ships them to the reward function on the host (``dataloader.py:119-139``: rows
    ``label_start_ix[v]:label_end_ix[v]`` of ``labels``).  Upload ``labels`` once,
    then a training step passes only its video indices: ``store.batch(ids)`` gathers
    the rows on the device into a :class:`RefSet`.
    """

    def __init__(self, labels, start_ix, end_ix, device=None):
        t = labels if torch.is_tensor(labels) else torch.from_numpy(_narrow(np.ascontiguousarray(labels, np.int64)))
        self.labels = t.to(device if device is not None else 'cuda').contiguous()
        self.start = np.ascontiguousarray(start_ix, dtype=np.int64)
        self.end = np.ascontiguousarray(end_ix, dtype=np.int64)
        if self.start.shape != self.end.shape or np.any(self.end < self.start):
            raise ValueError('start_ix / end_ix must pair up with end >= start')
        if len(self.end) and self.end.max() > self.labels.shape[0]:
            raise ValueError('end_ix beyond the label matrix')

    def batch(self, video_ids):
        ids = np.ascontiguousarray(video_ids, dtype=np.int64)
        counts = self.end[ids] - self.start[ids]
        offsets = np.concatenate([[0], np.cumsum(counts)])
        rows = np.repeat(self.start[ids] - offsets[:-1], counts) + np.arange(offsets[-1])
        picked = self.labels.index_select(0, torch.from_numpy(rows).to(self.labels.device))
        if len(counts) and np.all(counts == counts[0]):
            return RefSet(picked, per_video=int(counts[0]))
        return RefSet(picked, offsets=offsets)


class CookedRefs:
    """References staged once: the reference-side half of the scoring
    kernel -- canonicalisation, n-gram table with presence masks, norms, lengths -- is run one
    time per video by ``cst_refs_cook`` and its result kept in device memory, one image per
    video.  The ground-truth captions never change during training, yet the reference cooks
    them again on every call (``cook_refs`` under ``utils.py:208,274``).

    ``CookedRefs.cook(refs, scorer, use_eos)`` builds the images from a :class:`RefSet` /
    ``(V, R, W)`` tensor / list of arrays; ``cooked.batch(video_ids)`` selects a step's videos
    (only the indices move); the result is accepted wherever ``data_gts`` is.  Images are
    tied to the scorer's df table and to ``use_eos``.  An image is ONE table: at most 32 refs and
    1,100 n-gram instances per video.  Videos beyond that cannot be staged; ``cook`` finds them
    (``invalid``), and ``batch(ids)`` with host-side ids hands back the raw rows of a batch that
    names one of them, so the call scores it from raw rows instead of yielding NaN rewards.
    """

    def __init__(self, images, stride, ref_width, max_refs, use_eos, table, index=None, n_videos=None,
                 bad_rows=None, source=None, invalid=None):
        self.images, self.stride = images, int(stride)
        self.ref_width, self.max_refs, self.use_eos = int(ref_width), int(max_refs), int(bool(use_eos))
        self.table = table
        self.index = index
        self.n_videos = int(n_videos if n_videos is not None else images.shape[0])
        self.bad_rows = bad_rows            # device uint32[1]: out-of-range ref rows + invalid images
        self.source = source                # the raw rows, kept only when some video could not be staged
        self.invalid = invalid              # host bool per staged video, or None when all are valid

    @classmethod
    def cook(cls, data_gts, scorer, use_eos=0, verify=True):
        """``verify``: look the images over once they are written (one host sync; cooking is a one-off),
        warn about videos that could not be staged and keep their raw rows for ``batch``."""
        table = scorer.table
        dev = table.device
        refs = RefSet.coerce(data_gts, dev)
        if refs.max_refs > 32:
            raise ValueError('at most 32 refs per video can be staged (got %d)' % refs.max_refs)
        w = refs.rows.shape[1]
        max_refs = max(refs.max_refs, 1)
        stride = int(lib.cst_refs_cooked_bytes(w, max_refs))
        if stride < 0:
            raise ValueError('rows of width %d / %d refs per video cannot be staged' % (w, max_refs))
        images = torch.empty((refs.n_videos, stride), dtype=torch.uint8, device=dev)
        bad = torch.zeros(1, dtype=torch.int32, device=dev)
        check(lib.cst_refs_cook(table.handle, refs.rows.data_ptr(), _row_dtype(refs.rows), refs.rows.shape[0], w,
                                refs.rows.stride(0), refs.offsets.data_ptr() if refs.offsets is not None else None,
                                max_refs, refs.n_videos, int(use_eos), images.data_ptr(), stride, bad.data_ptr(),
                                _stream(dev)))
        source = invalid = None
        if verify and int(bad.item()):
            import warnings
            magic = images[:, :4].contiguous().view(torch.int32)[:, 0]
            invalid = (magic != 0x43535432).cpu().numpy()                   # "CST2": a valid image's header
            source = refs if invalid.any() else None
            warnings.warn('%d videos beyond one staged image (32 refs, 1,100 n-gram instances) are scored from raw '
                          'rows; %d reference rows hold ids outside [0, 65534]'
                          % (int(invalid.sum()), int(bad.item()) - int(invalid.sum())), RuntimeWarning, stacklevel=2)
            if not invalid.any():
                invalid = None
        return cls(images, stride, w, max_refs, use_eos, table, bad_rows=bad, source=source, invalid=invalid)

    def batch(self, video_ids):
        """The step's videos, in order: only the int64 indices go to the device."""
        if torch.is_tensor(video_ids):
            idx = video_ids.to(self.images.device, torch.int64).contiguous()
        else:
            ids = np.ascontiguousarray(video_ids, dtype=np.int64)
            if ids.size and (ids.min() < 0 or ids.max() >= self.images.shape[0]):
                raise ValueError('video index outside the %d staged videos' % self.images.shape[0])
            if self.invalid is not None and ids.size and self.invalid[ids].any():
                return self.source.take(ids)        # a video that could not be staged: this batch goes raw
            idx = torch.from_numpy(ids).to(self.images.device)
        return CookedRefs(self.images, self.stride, self.ref_width, self.max_refs, self.use_eos, self.table,
                          index=idx, n_videos=idx.numel(), bad_rows=self.bad_rows, source=self.source,
                          invalid=self.invalid)

    def nbytes(self):
        return self.images.numel()

    def __len__(self):
        return self.n_videos

    def __getitem__(self, videos):
        """A slice of this batch's videos (what ``dist.shard_batch`` takes of ``data_gts``)."""
        if not isinstance(videos, slice):
            raise TypeError('CookedRefs supports slices; use batch(ids) for arbitrary videos')
        lo, hi, step = videos.indices(self.n_videos)
        if self.index is not None:
            return self.batch(self.index[lo:hi:step])
        return self.batch(torch.arange(lo, hi, step, device=self.images.device))


def _narrow(rows):
    """Host int64 rows -> int32 when every id fits (halves the upload; ids outside
    [0, 65534] are rejected by the kernels either way)."""
    if rows.size and (rows.max() > np.iinfo(np.int32).max or rows.min() < np.iinfo(np.int32).min):
        return rows
    return rows.astype(np.int32)


def _rows_on_device(x, device):
    if torch.is_tensor(x):
        t = x.to(device)
    else:
        x = np.ascontiguousarray(x)
        if x.dtype == np.int64:
            x = _narrow(x)
        t = torch.from_numpy(x).to(device)
    if t.dtype not in (torch.int16, torch.int32, torch.int64):
        t = t.to(torch.int64)
    if t.dim() != 2:
        raise ValueError('caption rows must be 2-D (captions, tokens)')
    if t.stride(1) != 1:
        t = t.contiguous()
    return t


_scratch = {}


def _scratch_for(device, n_groups, stream=None):
    """Zero-initialised scratch the kernels leave clean (see cst_reward_scratch_bytes).
    One buffer per (device, stream): launches that may overlap never share one."""
    need = 256 + 16 * int(n_groups)                     # == cst_reward_scratch_bytes(n_groups)
    key = (device, stream if stream is not None else torch.cuda.current_stream(device).cuda_stream)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.zeros(max(need, 1 << 16), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def _launch_reward(scorer, sampled, greedy, refs, cands_per_video, use_eos, mode, seq_per_img, k,
                   bcmr, want_scores):
    dev = scorer.table.device
    n = sampled.shape[0]
    n_videos = n // cands_per_video
    if n_videos * cands_per_video != n or n_videos > refs.n_videos:
        raise ValueError('caption count %d does not match %d videos x %d' % (n, refs.n_videos, cands_per_video))
    reward = torch.empty(n, dtype=torch.float32, device=dev)
    stats = (torch.zeros if n == 0 else torch.empty)(_lib.N_STATS, dtype=torch.float64, device=dev)
    fused = cands_per_video == seq_per_img
    if n == 0:                                           # an empty shard: zero statistics, nothing to launch
        return reward, stats, (torch.empty(0, dtype=torch.float64, device=dev) if want_scores or not fused else None)
    scores = None
    if want_scores or not fused:
        scores = torch.empty(n * (2 if greedy is not None else 1), dtype=torch.float64, device=dev)
    stream = _stream(dev)                               # looked up once: it is 6 us of a 30 us call
    scratch = _scratch_for(dev, max(n_videos, n // seq_per_img), stream)
    if isinstance(refs, CookedRefs):
        if refs.table is not scorer.table or refs.use_eos != int(bool(use_eos)):
            raise ValueError('these references were staged for another df table or use_eos')
        check(lib.cst_reward_cooked(
            scorer.table.handle, sampled.data_ptr(), greedy.data_ptr() if greedy is not None else None,
            _row_dtype(sampled), n, sampled.shape[1], sampled.stride(0),
            refs.images.data_ptr(), refs.stride, refs.index.data_ptr() if refs.index is not None else None,
            refs.ref_width, refs.max_refs, n_videos, cands_per_video, int(use_eos), mode, seq_per_img, k,
            bcmr.data_ptr() if bcmr is not None else None,
            reward.data_ptr(), scores.data_ptr() if scores is not None else None, stats.data_ptr(),
            scratch.data_ptr(), scratch.numel(), stream))
        return reward, stats, scores
    check(lib.cst_reward(
        scorer.table.handle, sampled.data_ptr(), greedy.data_ptr() if greedy is not None else None,
        _row_dtype(sampled), n, sampled.shape[1], sampled.stride(0),
        refs.rows.data_ptr(), _row_dtype(refs.rows), refs.rows.shape[1], refs.rows.stride(0),
        refs.offsets.data_ptr() if refs.offsets is not None else None, refs.max_refs, n_videos,
        cands_per_video, int(use_eos), mode, seq_per_img, k,
        bcmr.data_ptr() if bcmr is not None else None,
        reward.data_ptr(), scores.data_ptr() if scores is not None else None, stats.data_ptr(),
        scratch.data_ptr(), scratch.numel(), stream))
    return reward, stats, scores


def _finish(reward, stats, width):
    rewards = reward.unsqueeze(1).expand(reward.shape[0], width)
    return rewards, stats[5], stats[6]          # the means are written by the kernel: no extra launch


def _need_scorer(bcmr_scorer):
    from .metrics import _RowScorer
    if not isinstance(bcmr_scorer, (CiderD, _RowScorer)):
        raise TypeError('the GPU reward path implements the CiderD, Bleu and Rouge scorers '
                        '(train.py:119-124; utils.py:194,260 dispatch); got %r' % type(bcmr_scorer).__name__)
    return bcmr_scorer


def _scorer_device(scorer):
    return scorer.table.device if isinstance(scorer, CiderD) else scorer.device


def _reward_from_scores(scorer, sampled, greedy, refs, cands_per_video, use_eos, mode, seq_per_img, k, bcmr):
    """Bleu / Rouge (train.py:119-124): their scores come from a kernel of their own; the baseline
    block of utils.py:219-224, 292-322 is then the standalone ``cst_scb_apply`` on the score vector
    (for ``Bleu`` the BLEU-4 list, utils.py:210-213, 276-279)."""
    if isinstance(refs, CookedRefs):
        raise TypeError('staged references hold the CIDEr-D half of the scoring: pass the rows (RefSet) to %s'
                        % type(scorer).__name__)
    dev = scorer.device
    n = sampled.shape[0]
    reward = torch.empty(n, dtype=torch.float32, device=dev)
    if n == 0:
        return reward, torch.zeros(_lib.N_STATS, dtype=torch.float64, device=dev)
    if n % seq_per_img:
        raise ValueError('cannot reshape %d scores into rows of %d' % (n, seq_per_img))
    stats = torch.empty(_lib.N_STATS, dtype=torch.float64, device=dev)
    bad = torch.zeros(1, dtype=torch.int32, device=dev)
    scores = scorer.score_rows(sampled, refs, cands_per_video, use_eos, greedy=greedy, bad=bad)
    other = scores[n:] if greedy is not None else bcmr
    stream = _stream(dev)
    scratch = _scratch_for(dev, n // seq_per_img, stream)
    check(lib.cst_scb_apply(scores.data_ptr(), other.data_ptr() if other is not None else None, n // seq_per_img,
                            seq_per_img, mode, k, reward.data_ptr(), stats.data_ptr(), scratch.data_ptr(),
                            scratch.numel(), dev.index, stream))
    stats[4:5] = bad                                    # rows with ids outside [-1, 65534], as cst_reward counts them
    return reward, stats


def get_self_critical_reward(model_res, greedy_res, data_gts, bcmr_scorer, expand_feat=0,
                             seq_per_img=20, use_eos=0, return_stats=False):
    """``utils.py:169-226``: reward = CIDEr-D(sampled) - CIDEr-D(greedy)."""
    scorer = _need_scorer(bcmr_scorer)
    dev = _scorer_device(scorer)
    sampled = _rows_on_device(model_res, dev)
    greedy = _rows_on_device(greedy_res, dev)
    if greedy.shape[0] != sampled.shape[0]:
        raise ValueError('greedy_res and model_res differ in caption count')
    if greedy.shape[1] != sampled.shape[1] or greedy.dtype != sampled.dtype or \
            greedy.stride(0) != sampled.stride(0):
        # the ABI takes one geometry for both row sets: pad the narrower one with <end>
        w = max(greedy.shape[1], sampled.shape[1])
        dt = torch.int64 if torch.int64 in (greedy.dtype, sampled.dtype) else torch.int32
        sampled = torch.nn.functional.pad(sampled.to(dt), (0, w - sampled.shape[1])).contiguous()
        greedy = torch.nn.functional.pad(greedy.to(dt), (0, w - greedy.shape[1])).contiguous()
    refs = RefSet.coerce(data_gts, dev)
    per_video = seq_per_img if expand_feat == 1 else 1
    if isinstance(scorer, CiderD):
        reward, stats, _ = _launch_reward(scorer, sampled, greedy, refs, per_video, use_eos,
                                          _lib.BASE_GREEDY, per_video, 0, None, False)
    else:
        reward, stats = _reward_from_scores(scorer, sampled, greedy, refs, per_video, use_eos, _lib.BASE_GREEDY,
                                            per_video, 0, None)
    out = _finish(reward, stats, model_res.shape[1])
    return out + (stats,) if return_stats else out


def get_cst_reward(model_res, data_gts, bcmr_scorer, bcmrscores=None, expand_feat=0, seq_per_img=20,
                   scb_captions=20, scb_baseline=1, use_eos=0, use_mixer=0, return_stats=False):
    """``utils.py:229-324``.  ``scb_baseline``: 1 = baseline from the GT consensus
    scores ``bcmrscores``, 2 = from the model samples' own scores; the baseline is
    the mean of the ``scb_captions`` lowest scores of the video's row."""
    scorer = _need_scorer(bcmr_scorer)
    dev = _scorer_device(scorer)
    width = model_res.shape[1]

    if scb_captions > 0:
        if scb_baseline == 1:
            mode = _lib.BASE_SCB_GT
            if bcmrscores is None:
                raise ValueError('bcmrscores is not set!')
        elif scb_baseline == 2:
            mode = _lib.BASE_SCB_SAMPLE
        else:
            raise ValueError('unknown scb_baseline!')
        k = min(int(scb_captions), seq_per_img)      # numpy's [:k] clamps the same way
    else:
        mode, k = _lib.BASE_NONE, 0

    bcmr = None
    if bcmrscores is not None:
        bcmr = bcmrscores if torch.is_tensor(bcmrscores) else torch.from_numpy(np.asarray(bcmrscores))
        bcmr = bcmr.to(dev, torch.float64).contiguous()

    if bcmrscores is None or use_mixer == 1:
        sampled = _rows_on_device(model_res, dev)
        if sampled.shape[0] % seq_per_img:
            raise ValueError('cannot reshape %d scores into rows of %d' % (sampled.shape[0], seq_per_img))
        refs = RefSet.coerce(data_gts, dev)
        per_video = seq_per_img if expand_feat == 1 else 1
        if bcmr is not None and mode == _lib.BASE_SCB_GT and bcmr.numel() != sampled.shape[0]:
            raise ValueError('bcmrscores must hold one score per sampled caption')
        gt = bcmr if mode == _lib.BASE_SCB_GT else None
        if isinstance(scorer, CiderD):
            reward, stats, _ = _launch_reward(scorer, sampled, None, refs, per_video, use_eos, mode,
                                              seq_per_img, k, gt, False)
        else:
            reward, stats = _reward_from_scores(scorer, sampled, None, refs, per_video, use_eos, mode,
                                                seq_per_img, k, gt)
    else:
        # WXE (utils.py:286-288): the precomputed consensus scores ARE the scores
        if bcmr.dim() != 2 or bcmr.shape[1] != seq_per_img:
            raise ValueError('bcmrscores must be (videos, seq_per_img)')
        n_rows = bcmr.shape[0]
        reward = torch.empty(n_rows * seq_per_img, dtype=torch.float32, device=dev)
        stats = torch.empty(_lib.N_STATS, dtype=torch.float64, device=dev)
        scratch = _scratch_for(dev, n_rows)
        check(lib.cst_scb_apply(bcmr.data_ptr(), bcmr.data_ptr(), n_rows, seq_per_img, mode, k,
                                reward.data_ptr(), stats.data_ptr(), scratch.data_ptr(),
                                scratch.numel(), dev.index, _stream(dev)))
    out = _finish(reward, stats, width)
    return out + (stats,) if return_stats else out
