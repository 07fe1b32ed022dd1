"""ORACLE (test infrastructure): float64 direct DFT of the measurement equation.

This is the definition that ``ducc0.wgridder.ms2dirty`` (reference call site
src/ska_sdp_cip/invert.py:170-183) approximates to relative L2 accuracy epsilon.
PARITY UNPINNED against ducc0 itself (not installable here).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .wstack import _p, lib


def dft_dirty(uvw, freq, vis, wgt, nx, ny, px, py, pixels=None):
    """Full image (nx,ny) float64, or values at flat pixel indices ``pixels``."""
    uvw = np.ascontiguousarray(uvw, dtype=np.float64)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    val = np.asarray(vis, dtype=np.complex128)
    if wgt is not None:
        val = val * np.asarray(wgt, dtype=np.float64)
    vr = np.ascontiguousarray(val.real)
    vi = np.ascontiguousarray(val.imag)
    if pixels is None:
        out = np.empty((nx, ny), dtype=np.float64)
        pp, n = None, 0
    else:
        pix = np.ascontiguousarray(pixels, dtype=np.int64)
        out = np.empty(pix.size, dtype=np.float64)
        pp, n = _p(pix), pix.size
    lib().oracle_dft_dirty(_p(uvw), C.c_int64(len(uvw)), _p(freq), C.c_int64(len(freq)), _p(vr), _p(vi),
                           C.c_int(nx), C.c_int(ny), C.c_double(px), C.c_double(py), pp, C.c_int64(n), _p(out))
    return out


def rel_rms(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.sqrt(np.sum((a - b) ** 2) / np.sum(b ** 2)))
