"""Device-resident document-frequency table.

Replaces the ``{'document_frequency': {ngram tuple: df}, 'ref_len': n}`` pickle
of ``compute_ciderdf.py:123-129`` as the scorer holds it after
``CiderD(df=path)`` (``train.py:121``): an open-addressing hash table in HBM
keyed by packed 1..4-gram token ids (see ``include/cst_reward.h``).
"""
from __future__ import annotations

import ctypes as C
import pickle
from collections import defaultdict

import numpy as np
import torch

from . import _lib
from ._lib import lib, check


def pack_ngrams(ngrams):
    """iterable of n-grams (tuples of ints or decimal strings, len 1..4) ->
    uint64 keys: sum (t_i + 1) << 16 i."""
    keys = np.zeros(len(ngrams), dtype=np.uint64)
    for i, g in enumerate(ngrams):
        k = 0
        if not 1 <= len(g) <= 4:
            raise ValueError('n-gram order must be 1..4, got %r' % (g,))
        for j, t in enumerate(g):
            t = int(t)
            if not 0 <= t <= _lib.MAX_TOKEN_ID:
                raise ValueError('token id %d outside [0, %d]' % (t, _lib.MAX_TOKEN_ID))
            k |= (t + 1) << (16 * j)
        keys[i] = k
    return keys


def unpack_keys(keys):
    """uint64 keys -> list of tuples of decimal strings (the pickle's key type,
    ``compute_ciderdf.py:84``)."""
    out = []
    for k in np.asarray(keys, dtype=np.uint64).tolist():
        g = []
        while k:
            g.append(str((k & 0xffff) - 1))
            k >>= 16
        out.append(tuple(g))
    return out


def idf_lut(n_documents):
    """``lut[d] = log(n_documents) - log(max(1, d))`` for d = 0..n_documents with the
    scalar numpy calls the scorer's ``counts2vec`` makes, so every value is
    bit-identical to it (``lut[0]`` is the idf of an unseen n-gram)."""
    ref_len = np.log(float(n_documents))
    return np.array([ref_len - np.log(max(1.0, float(d))) for d in range(int(n_documents) + 1)],
                    dtype=np.float64)


def _device_index(device):
    dev = torch.device(device if device is not None else 'cuda')
    if dev.type != 'cuda':
        raise ValueError('DfTable lives on a CUDA device, got %r' % (device,))
    return dev.index if dev.index is not None else torch.cuda.current_device()


class DfTable:
    """Handle of a device df table (``cst_df_table*``)."""

    def __init__(self, handle, n_documents):
        self._h = handle
        self.n_documents = int(n_documents)
        self.device = torch.device('cuda', lib.cst_df_device(handle))

    # -- construction ------------------------------------------------------
    @classmethod
    def from_arrays(cls, keys, df, n_documents, device=None):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        df = np.ascontiguousarray(df, dtype=np.float64)
        if keys.shape != df.shape:
            raise ValueError('keys and df differ in length')
        if len(df) and df.max() > n_documents:
            raise ValueError('document frequency above the document count')
        lut = idf_lut(n_documents)
        h = C.c_void_p()
        check(lib.cst_df_create(keys.ctypes.data, df.ctypes.data, len(keys), lut.ctypes.data,
                                int(n_documents), _device_index(device), C.byref(h)))
        return cls(h, n_documents)

    @classmethod
    def from_dict(cls, document_frequency, ref_len, device=None):
        """From the reference's mapping ``{tuple[str]: float}`` and document count
        (the two fields of the pickle, ``compute_ciderdf.py:123-129``)."""
        grams = list(document_frequency.keys())
        vals = np.fromiter((document_frequency[g] for g in grams), dtype=np.float64, count=len(grams))
        return cls.from_arrays(pack_ngrams(grams), vals, ref_len, device)

    @classmethod
    def from_pickle(cls, path, device=None):
        """Load a ``*_ciderdf.pkl`` written by the reference's compute_ciderdf."""
        with open(path, 'rb') as fh:
            blob = pickle.load(fh, encoding='latin1')
        return cls.from_dict(blob['document_frequency'], blob['ref_len'], device)

    @classmethod
    def from_rows(cls, rows, video_offsets=None, use_eos=0):
        """Build on the device from reference rows (``compute_ciderdf.py:56-93``;
        corpus mode of ``compute_scores.py:92``).  ``rows``: CUDA int tensor
        ``(n_videos, R, W)`` or ``(n_rows, W)`` with ``video_offsets`` (n_videos+1).
        Rows are canonicalised like any caption row; to reproduce the reference's
        df pickle pass rows holding the words followed by ``<end>`` with
        ``use_eos=1`` (``compute_ciderdf.py:115-118``)."""
        if not rows.is_cuda:
            raise ValueError('rows must be a CUDA tensor')
        if rows.dim() == 3:
            n_videos, r, w = rows.shape
            video_offsets = np.arange(n_videos + 1, dtype=np.int64) * r
            rows = rows.reshape(n_videos * r, w)
        elif video_offsets is None:
            raise ValueError('video_offsets is required for 2-D rows')
        rows = rows.contiguous()
        # the build runs on the legacy default stream and is synchronous: whatever produced `rows` on
        # the caller's (possibly non-blocking) stream has to be done first
        torch.cuda.current_stream(rows.device).synchronize()
        dtype = _row_dtype(rows)
        off = np.ascontiguousarray(video_offsets, dtype=np.int64)
        n_videos = len(off) - 1
        lut = idf_lut(n_videos)
        h = C.c_void_p()
        check(lib.cst_df_build_from_rows(rows.data_ptr(), dtype, rows.shape[0], rows.shape[1],
                                         rows.stride(0), off.ctypes.data, n_videos, int(use_eos),
                                         lut.ctypes.data, rows.device.index, C.byref(h)))
        return cls(h, n_videos)

    # -- access --------------------------------------------------------------
    def __len__(self):
        return int(lib.cst_df_size(self._h))

    @property
    def capacity(self):
        return int(lib.cst_df_capacity(self._h))

    @property
    def handle(self):
        if self._h is None:
            raise ValueError('DfTable is closed')
        return self._h

    def lookup(self, keys):
        """-> (df, idf) float64 CUDA tensors for a uint64/int64 key tensor or array."""
        if not torch.is_tensor(keys):
            keys = torch.from_numpy(np.ascontiguousarray(keys, dtype=np.uint64).view(np.int64))
        keys = keys.to(self.device).contiguous()
        n = keys.numel()
        df = torch.empty(n, dtype=torch.float64, device=self.device)
        idf = torch.empty(n, dtype=torch.float64, device=self.device)
        check(lib.cst_df_lookup(self.handle, keys.data_ptr(), n, df.data_ptr(), idf.data_ptr(),
                                torch.cuda.current_stream(self.device).cuda_stream))
        return df, idf

    def export(self):
        """-> (keys uint64, df float64) numpy arrays of every stored n-gram."""
        n = len(self)
        keys = np.empty(n, dtype=np.uint64)
        df = np.empty(n, dtype=np.float64)
        check(lib.cst_df_export(self.handle, keys.ctypes.data, df.ctypes.data))
        return keys, df

    def max_token_id(self):
        """Largest token id any stored n-gram holds (-1 for an empty table); computed once."""
        if getattr(self, '_max_token', None) is None:
            keys, _ = self.export()
            m = 0
            for shift in (0, 16, 32, 48):
                if len(keys):
                    m = max(m, int(((keys >> np.uint64(shift)) & np.uint64(0xffff)).max()))
            self._max_token = m - 1
        return self._max_token

    def to_pickle(self, path):
        """Write the table in the reference's schema (``compute_ciderdf.py:123-129``)
        so the reference's own scorer can consume a device-built table."""
        keys, df = self.export()
        table = defaultdict(float)
        for g, v in zip(unpack_keys(keys), df.tolist()):
            table[g] = v
        with open(path, 'wb') as fh:
            pickle.dump({'document_frequency': table, 'ref_len': self.n_documents}, fh,
                        protocol=pickle.HIGHEST_PROTOCOL)

    def close(self):
        if self._h is not None:
            lib.cst_df_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _row_dtype(t):
    if t.dtype == torch.int32:
        return _lib.DTYPE_I32
    if t.dtype == torch.int64:
        return _lib.DTYPE_I64
    if t.dtype == torch.int16:
        return _lib.DTYPE_I16
    raise ValueError('token rows must be int16, int32 or int64, got %s' % t.dtype)
