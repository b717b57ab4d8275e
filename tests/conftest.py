import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')
    # the package refuses to import without its CUDA library (no CPU fallback): build it
    # (nvcc cross-compiles without a GPU) before any test module imports the package
    import __graft_entry__ as graft
    graft.build()


def pytest_collection_modifyitems(config, items):
    # GPU tests are selected with -m gpu; when no device is present they are
    # skipped (never silently passed through some CPU path).
    try:
        import torch
        have = torch.cuda.is_available()
    except Exception:
        have = False
    if have:
        return
    skip = pytest.mark.skip(reason='no CUDA device in this container')
    for item in items:
        if 'gpu' in item.keywords:
            item.add_marker(skip)


def ids_to_df_dict(ngram_ids, df):
    """fixture arrays -> the reference's dict{tuple[str]: float}."""
    out = {}
    for row, val in zip(ngram_ids.tolist(), df.tolist()):
        out[tuple(str(t) for t in row if t >= 0)] = val
    return out


@pytest.fixture(scope='session')
def golden_cases():
    return np.load(os.path.join(GOLDEN, 'reward_cases.npz'))


@pytest.fixture(scope='session')
def golden_df_v10k():
    z = np.load(os.path.join(GOLDEN, 'df_v10k.npz'))
    return z, ids_to_df_dict(z['ngram_ids'], z['df']), int(z['ref_len'])


@pytest.fixture(scope='session')
def golden_df_v40():
    z = np.load(os.path.join(GOLDEN, 'df_v40.npz'))
    return z, ids_to_df_dict(z['ngram_ids'], z['df']), int(z['ref_len'])
