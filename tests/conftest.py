import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
ORACLE_SO = os.path.join(ROOT, "oracle", "libdoma_oracle.so")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: minutes on CPU")


def _build_oracle():
    import subprocess

    src = os.path.join(ROOT, "oracle", "doma_oracle.c")
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, capture_output=True)


@pytest.fixture(scope="session")
def oracle_lib():
    """The CPU oracle (test infrastructure) behind the same ctypes binding as the product library."""
    from doma_b200 import abi

    _build_oracle()
    return abi.Library(ORACLE_SO, "doma_oracle_")


@pytest.fixture(scope="session")
def cuda_lib():
    """The product library.  GPU tests fail (not skip) when it is missing: there is no fallback."""
    from doma_b200 import abi

    return abi.load_library()


def golden_names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN) if f.endswith(".npz"))


def load_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: g[k] for k in g.files}


def golden_case_path(name):
    return os.path.join(GOLDEN, "cases", name + ".xml")


def rel_err(a, b):
    """max |a-b| / max(|b|) per group-array: the flux tolerance of north_star is relative."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    scale = np.maximum(np.abs(b), 1e-300)
    return float(np.max(np.abs(a - b) / scale)) if a.size else 0.0
