"""Shared test configuration.

`-m "not gpu"`: oracle vs golden vectors, host logic, C-ABI symbol check, gloo multi-process logic.
`-m gpu`: parity tests proper -- the CUDA path, called through the C-ABI, against the oracle.
"""
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "ska-sdp-continuum-imaging-pipeline_b200"
for p in (str(ROOT), str(PKG)):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    def load(name):
        with np.load(GOLDEN / f"{name}.npz") as z:
            return {k: z[k] for k in z.files}
    return load


@pytest.fixture(scope="session")
def built_lib():
    """libcip_b200.so, built on demand (nvcc cross-compiles without a GPU)."""
    from ska_sdp_cip import _cabi
    if not _cabi.lib_path().exists():
        import __graft_entry__ as g
        g.build()
    return _cabi.lib()


def random_problem(nrow=2000, nchan=8, nx=128, fov_scale=1.0, wscale=60.0, seed=1, uvscale=800.0, flag=0.05):
    """Small random invert problem (uvw m, freq Hz, vis c64, wgt f32, npix, pixel size)."""
    rng = np.random.default_rng(seed)
    uvw = rng.normal(size=(nrow, 3)) * np.array([uvscale, uvscale, wscale])
    freq = 1.3e9 + 1e6 * np.arange(nchan)
    vis = (rng.normal(size=(nrow, nchan)) + 1j * rng.normal(size=(nrow, nchan))).astype(np.complex64)
    wgt = rng.uniform(0.5, 1.5, size=(nrow, nchan)).astype(np.float32)
    wgt[rng.random(wgt.shape) < flag] = 0
    px = fov_scale * 0.4 / (np.abs(uvw[:, :2]).max() * freq.max() / 299792458.0)
    return uvw, freq, vis, wgt, nx, px
