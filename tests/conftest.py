import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def random_cloud(n, extent, seed, batch=1, lo=0):
    """n unique random voxels in [lo, lo+extent)^3 per batch element -> [N,4] int32 (b,x,y,z)."""
    rng = np.random.default_rng(seed)
    out = []
    for b in range(batch):
        p = np.unique(rng.integers(lo, lo + extent, size=(n, 3)), axis=0)
        out.append(np.concatenate([np.full((p.shape[0], 1), b), p], axis=1))
    c = np.concatenate(out, 0).astype(np.int32)
    return c[rng.permutation(c.shape[0])]


@pytest.fixture(scope="session")
def pcc():
    import pcc_b200
    return pcc_b200


@pytest.fixture(scope="session")
def oracle():
    import oracle as o
    o.prims.build()
    return o
