"""Shared fixtures.  `-m "not gpu"`: oracle vs golden vectors, host logic, ABI
surface.  `-m gpu`: the parity tests proper, through the C ABI on a B200.

Nothing here (or in any test) reads /root/reference at run time: the golden
vectors under tests/golden/ were copied from it by tools/make_golden.py.
"""
import gzip
import hashlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def gold_path(name):
    p = os.path.join(GOLD, name)
    return p if os.path.exists(p) else p + ".gz"


def load_instance(sample):
    from recommender_system_b200.io import read_instance
    return read_instance(gold_path(sample + ".in"))


def load_out(sample):
    from recommender_system_b200.io import read_golden_out
    return read_golden_out(gold_path(sample + ".out"))


def lr_sha256(L, R):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(L).tobytes())
    h.update(np.ascontiguousarray(R).tobytes())
    return h.hexdigest()


@pytest.fixture(scope="session")
def expected():
    with open(os.path.join(GOLD, "expected.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def oracle():
    """The CPU checker (oracle/), compiled on demand.  Test infrastructure only."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    pyoracle.build(ref=False)
    pyoracle.lib()
    return pyoracle


@pytest.fixture(scope="session")
def built_lib():
    """libmatfact_b200.so, built by `make` if missing (nvcc cross-compiles without a GPU)."""
    import subprocess
    from recommender_system_b200 import api
    if not os.path.exists(api.LIB_PATH) or not os.path.exists(os.path.join(ROOT, "bin", "matFact")):
        subprocess.run(["make", "-C", ROOT, "-j4"], check=True, stdout=subprocess.DEVNULL)
    return api


@pytest.fixture(scope="session")
def gpu(built_lib):
    api = built_lib
    if api.device_count() < 1:
        pytest.fail("no B200 visible: the -m gpu tests must run on the GPU box")
    return api
