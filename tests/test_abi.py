"""Host-side checks that need no GPU: the C-ABI library loads, exports every
symbol include/cst_reward.h declares, and refuses to compute without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import __graft_entry__ as graft

from conftest import ROOT


@pytest.fixture(scope='module')
def cst():
    graft.build()
    import cst_captioning_b200
    return cst_captioning_b200


def _declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'cst_reward.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(cst_[a-z0-9_]+)\s*\(', text)))


def test_header_symbols_are_exported(cst):
    names = _declared_symbols()
    assert len(names) >= 14
    raw = C.CDLL(cst.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), 'libcst_reward.so lacks %s' % n
    from cst_captioning_b200 import _lib
    assert sorted(_lib.SYMBOLS) == names          # the binding covers exactly the header
    assert cst.lib.cst_abi_version() == 1


def test_every_entry_point_is_mapped_to_the_reference_in_the_docs():
    """INTEGRATION.md names, for each C symbol, the reference interface it replaces; the header
    cites reference file:line somewhere in the comment above each group of declarations."""
    doc = open(os.path.join(ROOT, 'INTEGRATION.md')).read()
    for n in _declared_symbols():
        assert '`%s`' % n in doc, '%s is missing from INTEGRATION.md' % n
    header = open(os.path.join(ROOT, 'include', 'cst_reward.h')).read()
    for cited in ('utils.py:', 'train.py:', 'compute_ciderdf.py:', 'model.py:', 'dataloader.py:'):
        assert cited in header, 'the header no longer cites %s' % cited


def test_scratch_bytes_and_constants(cst):
    assert cst.lib.cst_reward_scratch_bytes(0) == 256
    assert cst.lib.cst_reward_scratch_bytes(65536) == 256 + 16 * 65536
    assert cst.lib.cst_reward_scratch_bytes(-1) == -1


def test_pack_unpack_roundtrip(cst):
    grams = [('0',), ('5', '0'), ('3', '4', '5'), ('65534', '0', '7', '9'), (12, 13)]
    keys = cst.pack_ngrams(grams)
    assert keys[0] == 1 and keys[1] == 6 | (1 << 16)
    assert cst.unpack_keys(keys) == [tuple(str(t) for t in g) for g in grams]
    assert len(set(keys.tolist())) == len(grams)
    with pytest.raises(ValueError):
        cst.pack_ngrams([('65535',)])
    with pytest.raises(ValueError):
        cst.pack_ngrams([('1', '2', '3', '4', '5')])


def test_idf_lut_matches_scorer_formula(cst):
    from cst_captioning_b200.df_table import idf_lut
    lut = idf_lut(60)
    assert len(lut) == 61
    for d in range(61):
        assert lut[d] == np.log(float(60)) - np.log(max(1.0, float(d)))     # bit-identical
    assert lut[60] == 0.0 and lut[0] == np.log(60.0)


def test_no_silent_cpu_fallback(cst):
    """Without a CUDA device every compute entry point must fail loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    keys = cst.pack_ngrams([('3',), ('3', '4')])
    with pytest.raises((cst.CstError, ValueError, RuntimeError)):
        cst.DfTable.from_arrays(keys, np.array([1.0, 1.0]), 5, device='cuda:0')


def test_argument_validation_in_c(cst):
    lib = cst.lib
    out = C.c_void_p()
    rc = lib.cst_df_create(None, None, 3, None, 5, 0, C.byref(out))
    assert rc == -1 and b'NULL' in lib.cst_last_error()
    rc = lib.cst_scb_apply(None, None, 4, 20, 9, 5, None, None, None, 0, 0, None)
    assert rc == -1 and b'unknown baseline mode' in lib.cst_last_error()
    rc = lib.cst_scb_apply(None, None, 4, 20, 3, 21, None, None, None, 0, 0, None)
    assert rc == -1 and b'scb_k' in lib.cst_last_error()
    rc = lib.cst_ngram_counts(C.c_void_p(8), 0, 1, 40, 40, 0, C.c_void_p(8), C.c_void_p(8),
                              C.c_void_p(8), C.c_void_p(8), 0, None)
    assert rc == -4 and b'row width' in lib.cst_last_error()


def test_library_has_sm100a_code(cst):
    import subprocess
    out = subprocess.run(['cuobjdump', '-lelf', cst.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip('cuobjdump unavailable')
    assert 'sm_100a' in out.stdout
