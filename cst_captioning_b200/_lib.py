"""ctypes binding of ``libcst_reward.so`` (C ABI: ``include/cst_reward.h``).

There is no CPU fallback: if the shared library has not been built the import
fails loudly, and every compute call fails if no CUDA device is present.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CST_REWARD_LIB: developer override to load an experimental build of the same ABI
LIB_PATH = os.environ.get('CST_REWARD_LIB') or os.path.join(_HERE, 'libcst_reward.so')

OK, E_ARG, E_CUDA, E_NOMEM, E_RANGE = 0, -1, -2, -3, -4
DTYPE_I32, DTYPE_I64, DTYPE_I16 = 0, 1, 2
BASE_NONE, BASE_GREEDY, BASE_SCB_GT, BASE_SCB_SAMPLE = 0, 1, 2, 3
MAX_ROW_WIDTH = 32
MAX_TOKEN_ID = 65534
NGRAM_SLOTS_PER_ROW = 128
N_STATS = 7

# every symbol include/cst_reward.h declares: name -> (restype, argtypes)
_vp, _i32, _i64, _f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
_ROWS = [_vp, _vp, _i32, _i64, _i32, _i64,      # cand, greedy, dtype, n_cand, width, stride
         _vp, _i32, _i32, _i64, _vp,            # refs, dtype, width, stride, ref_offsets
         _i32, _i64, _i32, _i32]                # refs_per_video, n_videos, cands_per_video, use_eos
SYMBOLS = {
    'cst_last_error': (C.c_char_p, []),
    'cst_abi_version': (C.c_int, []),
    'cst_kernel_launches': (_i64, []),
    'cst_df_create': (C.c_int, [_vp, _vp, _i64, _vp, _i64, C.c_int, C.POINTER(_vp)]),
    'cst_df_build_from_rows': (C.c_int, [_vp, _i32, _i64, _i32, _i64, _vp, _i64, _i32, _vp, C.c_int,
                                         C.POINTER(_vp)]),
    'cst_df_destroy': (C.c_int, [_vp]),
    'cst_df_size': (_i64, [_vp]),
    'cst_df_capacity': (_i64, [_vp]),
    'cst_df_device': (C.c_int, [_vp]),
    'cst_df_export': (C.c_int, [_vp, _vp, _vp]),
    'cst_df_lookup': (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp]),
    'cst_ngram_counts': (C.c_int, [_vp, _i32, _i64, _i32, _i64, _i32, _vp, _vp, _vp, _vp, C.c_int, _vp]),
    'cst_cider_scores': (C.c_int, [_vp] + _ROWS + [_vp, _vp]),
    'cst_reward_scratch_bytes': (_i64, [_i64]),
    'cst_reward': (C.c_int, [_vp] + _ROWS + [_i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _i64, _vp]),
    'cst_reward_noalloc': (C.c_int, [_vp] + _ROWS + [_i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp]),
    'cst_refs_cooked_bytes': (_i64, [_i32, _i32]),
    'cst_refs_cook': (C.c_int, [_vp, _vp, _i32, _i64, _i32, _i64, _vp, _i32, _i64, _i32, _vp, _i64, _vp, _vp]),
    'cst_reward_cooked': (C.c_int, [_vp, _vp, _vp, _i32, _i64, _i32, _i64,      # table, cand, greedy, dtype, n, width, stride
                                    _vp, _i64, _vp, _i32, _i32, _i64,            # images, stride, index, ref_width, refs/video, n_videos
                                    _i32, _i32, _i32, _i32, _i32, _vp,           # cands/video, use_eos, mode, seq_per_img, scb_k, bcmr
                                    _vp, _vp, _vp, _vp, _i64, _vp]),             # reward, scores, stats, scratch, bytes, stream
    'cst_criterion_scratch_bytes': (_i64, []),
    'cst_reward_criterion_fwd': (C.c_int, [_vp, _i32, _i64, _vp, _i64, _vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _i64,
                                           C.c_int, _vp]),
    'cst_reward_criterion_bwd': (C.c_int, [_vp, _i32, _i64, _vp, _i64, _i64, _vp, _vp, _i64, _i32, _vp, _i64, C.c_int,
                                           _vp]),
    'cst_df_sole_holders': (C.c_int, [_vp, _vp, _i32, _i64, _i32, _i64, _i32, _i64, _i32, _vp, _vp, _vp]),
    'cst_df_leave_one_out': (C.c_int, [_vp, _vp, _i32, _vp]),
    'cst_bleu_scores': (C.c_int, _ROWS + [_vp, _vp, _vp, _i32, _vp]),
    'cst_rouge_scores': (C.c_int, _ROWS + [_vp, _vp, _i32, _vp]),
    'cst_scb_apply': (C.c_int, [_vp, _vp, _i64, _i32, _i32, _i32, _vp, _vp, _vp, _i64, C.c_int, _vp]),
}


class CstError(RuntimeError):
    """A call into libcst_reward.so failed (code, message from cst_last_error)."""

    def __init__(self, code, message):
        super().__init__('libcst_reward error %d: %s' % (code, message))
        self.code = code


def _load():
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            'cst_captioning_b200: %s is missing. Build it first with '
            '`python -c "import __graft_entry__ as g; g.build()"` (nvcc, sm_100a). '
            'There is no CPU fallback for the reward path.' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc):
    if rc != OK:
        msg = lib.cst_last_error().decode('utf-8', 'replace')
        if rc == E_ARG:
            raise ValueError('libcst_reward: ' + msg)
        raise CstError(rc, msg)
    return rc
