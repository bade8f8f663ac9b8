"""Shared fixtures. GPU tests are marked ``@pytest.mark.gpu``; everything else runs on a CPU-only machine."""
import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_names():
    return sorted(os.path.splitext(os.path.basename(f))[0] for f in glob.glob(os.path.join(GOLDEN, "*.npz")))


class Golden:
    """One committed fixture produced by tests/golden/make_golden.py from the reference's PyTorch code."""

    def __init__(self, name):
        import fvsrn_b200  # noqa: F401  (loader shim)
        from fvsrn_b200 import volnet
        self.name = name
        d = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.volnet_bytes = d["volnet"].tobytes()
        self.net = volnet.loads(self.volnet_bytes)
        self.positions = d["positions"]
        self.world = d["world"]
        self.screen = d["screen"]
        self.time = float(d["time"])
        self.has_dvr = "dvr" in d
        if self.has_dvr:
            self.dvr = d["dvr"]
            self.dvr_size = int(d["dvr_size"])
            self.dvr_step = float(d["dvr_step"])
            self.camera = d["camera"]


@pytest.fixture(scope="session")
def plume_tf():
    return np.load(os.path.join(GOLDEN, "plume_tf.npy"))


@pytest.fixture(scope="session")
def built_lib():
    """The product library must exist (built by __graft_entry__.build()); build it if the tree is fresh."""
    from fvsrn_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib.lib()


def psnr(a, b):
    mse = float(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))
    return 200.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)
