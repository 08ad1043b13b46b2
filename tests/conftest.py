import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return int(z["n"]), z["row_offsets"].astype(np.int64), z["col"].astype(np.int32), z["vals"]


@pytest.fixture(scope="session")
def shared_lib():
    return json.load(open(os.path.join(GOLDEN, "shared_lib.json")))


@pytest.fixture(scope="session")
def orc():
    import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def gr():
    """the product package; GPU tests fail loudly if the native library is absent"""
    import gunrockinst_b200 as g
    g._capi.lib()
    assert g.device_count() > 0, "no CUDA device visible"
    return g


_RMAT_CACHE = {}


def rmat_graph(orc, scale, edgefactor=16, undirected=True):
    key = (scale, edgefactor, undirected)
    if key not in _RMAT_CACHE:
        _RMAT_CACHE[key] = orc.rmat_csr(scale, edgefactor, undirected=undirected)
    return _RMAT_CACHE[key]
