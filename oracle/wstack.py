"""ORACLE / CPU BASELINE PORT driver (test infrastructure, never on the product path).

Drives oracle/csrc/wgrid.c (G1 scan + binning, G2 gridding, G4 w-screen and
correction) and uses scipy.fft (pocketfft, the FFT ducc0 itself ships) for G3.
Restates what the reference obtains from ``ducc0.wgridder.ms2dirty`` at
src/ska_sdp_cip/invert.py:170-183.  PARITY UNPINNED against ducc0 (absent here);
validated against oracle.dft to the kernel's design error.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import scipy.fft
from numpy.polynomial.legendre import leggauss

from . import build as _build
from .plan import Plan, make_plan

_lib = None


class _CPlan(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("nu", C.c_int), ("nv", C.c_int), ("W", C.c_int),
                ("nplanes", C.c_int), ("px", C.c_double), ("py", C.c_double), ("sigma", C.c_double),
                ("beta", C.c_double), ("dw", C.c_double), ("w0", C.c_double), ("nshift", C.c_double)]


class _Rec(C.Structure):
    _fields_ = [("i0", C.c_int32), ("j0", C.c_int32), ("p0", C.c_int32), ("fu", C.c_double), ("fv", C.c_double),
                ("fw", C.c_double), ("vr", C.c_double), ("vi", C.c_double)]


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(_build.build()))
        _lib.oracle_wg_bin.restype = C.c_int64
        _lib.oracle_num_threads.restype = C.c_int
        _lib.oracle_set_num_threads.argtypes = [C.c_int]
        _lib.oracle_set_num_threads(os.cpu_count() or 1)  # all host cores, whatever OMP_NUM_THREADS says
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _cplan(pl: Plan) -> _CPlan:
    return _CPlan(pl.nx, pl.ny, pl.nu, pl.nv, pl.W, pl.nplanes, pl.px, pl.py, pl.sigma, pl.beta, pl.dw, pl.w0,
                  pl.nshift)


def scan(uvw, freq, vis, wgt):
    L = lib()
    wmin, wmax, sumw, na = C.c_double(), C.c_double(), C.c_double(), C.c_int64()
    L.oracle_wg_scan(_p(uvw), C.c_int64(len(uvw)), _p(freq), C.c_int64(len(freq)), _p(vis),
                     _p(wgt) if wgt is not None else None, C.byref(wmin), C.byref(wmax), C.byref(na), C.byref(sumw))
    return wmin.value, wmax.value, na.value, sumw.value


def bin_records(uvw, freq, vis, wgt, pl: Plan):
    """The oracle's gridder-ready records of all ACTIVE samples (oracle_wg_bin, csrc/wgrid.c): structured array with
    i0, j0 (first cell, shifted coordinates), p0 (first plane), fu, fv, fw (offsets in [0, 1)), vr, vi."""
    L = lib()
    uvw = np.ascontiguousarray(uvw, dtype=np.float64)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    vis = np.ascontiguousarray(vis, dtype=np.complex64)
    if wgt is not None:
        wgt = np.ascontiguousarray(wgt, dtype=np.float32)
    nrow, nchan = vis.shape
    cp = _cplan(pl)
    ntu, ntv = (pl.nu + 32) // 32 + 1, (pl.nv + 32) // 32 + 1
    recs = (_Rec * (nrow * nchan))()
    offsets = np.zeros(ntu * ntv + 1, dtype=np.int64)
    n = L.oracle_wg_bin(C.byref(cp), _p(uvw), C.c_int64(nrow), _p(freq), C.c_int64(nchan), _p(vis),
                        _p(wgt) if wgt is not None else None, recs, _p(offsets), C.c_int(ntu), C.c_int(ntv))
    return np.frombuffer(recs, dtype=np.dtype(_Rec), count=n).copy()


def kernel_ft(pl: Plan, xi, nq=None):
    """c(xi) for the plan's ES kernel."""
    nq = nq or (2 * pl.W + 40)
    x, w = leggauss(nq)
    xi = np.ascontiguousarray(xi, dtype=np.float64)
    out = np.empty_like(xi)
    cp = _cplan(pl)
    lib().oracle_wg_chat(C.byref(cp), _p(x), _p(w), C.c_int(nq), _p(xi), C.c_int64(xi.size), _p(out))
    return out


def ms2dirty(uvw, freq, vis, wgt, nx, ny, px, py, epsilon=1e-4, *, fp64=False, sigma=None, W=None,
             plan: Plan | None = None, return_plan=False, planes_out=None):
    """CPU w-stacking dirty image, float64 result (nx, ny).

    ``fp64`` selects double grids (the oracle setting); False mirrors ms2dirty's
    single-precision path for complex64 input and is what the CPU baseline times.
    """
    L = lib()
    uvw = np.ascontiguousarray(uvw, dtype=np.float64)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    vis = np.ascontiguousarray(vis, dtype=np.complex64)
    if wgt is not None:
        wgt = np.ascontiguousarray(wgt, dtype=np.float32)
    nrow, nchan = vis.shape
    wmin, wmax, nact, _ = scan(uvw, freq, vis, wgt)
    img = np.zeros((nx, ny), dtype=np.float64)
    if nact == 0:
        return (img, None) if return_plan else img
    pl = plan or make_plan(nx, ny, px, py, epsilon, wmin, wmax, nact, fp64=fp64, sigma=sigma, W=W)
    cp = _cplan(pl)
    TS = 32
    ntu, ntv = (pl.nu + 32) // TS + 1, (pl.nv + 32) // TS + 1
    recs = (_Rec * (nrow * nchan))()
    offsets = np.zeros(ntu * ntv + 1, dtype=np.int64)
    L.oracle_wg_bin(C.byref(cp), _p(uvw), C.c_int64(nrow), _p(freq), C.c_int64(nchan), _p(vis),
                    _p(wgt) if wgt is not None else None, recs, _p(offsets), C.c_int(ntu), C.c_int(ntv))
    rdt, cdt = (np.float64, np.complex128) if fp64 else (np.float32, np.complex64)
    grid_fn = L.oracle_wg_grid_plane_f64 if fp64 else L.oracle_wg_grid_plane_f32
    scr_fn = L.oracle_wg_wscreen_f64 if fp64 else L.oracle_wg_wscreen_f32
    nthreads = os.cpu_count() or 1
    grid = np.empty((pl.nu, pl.nv), dtype=cdt)
    for p in range(pl.nplanes):
        grid.fill(0)
        grid_fn(C.byref(cp), recs, _p(offsets), C.c_int(ntu), C.c_int(ntv), C.c_int(p), _p(grid))
        if planes_out is not None:
            planes_out.append(grid.copy())
        F = scipy.fft.ifft2(grid, norm="forward", workers=nthreads)
        F = np.ascontiguousarray(F, dtype=cdt)
        scr_fn(C.byref(cp), _p(F), C.c_int(p), _p(img))
    nq = 2 * pl.W + 40
    x, w = leggauss(nq)
    cu = kernel_ft(pl, (np.arange(nx) - nx // 2) / pl.nu)
    cv = kernel_ft(pl, (np.arange(ny) - ny // 2) / pl.nv)
    L.oracle_wg_correct(C.byref(cp), _p(x), _p(w), C.c_int(nq), _p(cu), _p(cv), _p(img))
    return (img, pl) if return_plan else img


class _FastRec(C.Structure):
    _fields_ = [("i0", C.c_int32), ("j0", C.c_int32), ("p0", C.c_int32), ("fu", C.c_float), ("fv", C.c_float),
                ("fw", C.c_float), ("vr", C.c_float), ("vi", C.c_float)]


def ms2dirty_fast(uvw, freq, vis, wgt, nx, ny, px, py, epsilon=1e-4, *, sigma=None, W=None, max_stack_bytes=12 << 30):
    """The timed CPU baseline (csrc/wgrid_fast.c): float32 grids like ms2dirty's complex64 path, parallel binning,
    kernel taps evaluated once per record for all its planes (planes resident in blocks of ``max_stack_bytes``),
    pocketfft (scipy.fft) with all threads per plane, tabulated w correction.  Returns the float64 image."""
    L = lib()
    L.oracle_fast_bin.restype = C.c_int64
    uvw = np.ascontiguousarray(uvw, dtype=np.float64)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    vis = np.ascontiguousarray(vis, dtype=np.complex64)
    if wgt is not None:
        wgt = np.ascontiguousarray(wgt, dtype=np.float32)
    nrow, nchan = vis.shape
    wmin, wmax, nact, _ = scan(uvw, freq, vis, wgt)
    img = np.zeros((nx, ny), dtype=np.float64)
    if nact == 0:
        return img
    pl = make_plan(nx, ny, px, py, epsilon, wmin, wmax, nact, fp64=False, sigma=sigma, W=W)
    cp = _cplan(pl)
    ts = L.oracle_fast_tile_edge()
    ntu, ntv = (pl.nu + ts) // ts + 1, (pl.nv + ts) // ts + 1  # the last two tile rows / columns are the wrap-around seam
    recs = np.empty(nrow * nchan, dtype=np.dtype(_FastRec))
    offsets = np.zeros(ntu * ntv + 1, dtype=np.int64)
    L.oracle_fast_bin(C.byref(cp), _p(uvw), C.c_int64(nrow), _p(freq), C.c_int64(nchan), _p(vis),
                      _p(wgt) if wgt is not None else None, _p(recs), _p(offsets), C.c_int(ntu), C.c_int(ntv))
    nthreads = os.cpu_count() or 1
    plane_bytes = pl.nu * pl.nv * 8
    block = max(1, min(pl.nplanes, int(max_stack_bytes // plane_bytes)))
    for p_lo in range(0, pl.nplanes, block):
        p_hi = min(pl.nplanes, p_lo + block)
        stack = np.zeros((p_hi - p_lo, pl.nu, pl.nv), dtype=np.complex64)
        L.oracle_fast_grid(C.byref(cp), _p(recs), _p(offsets), C.c_int(ntu), C.c_int(ntv), C.c_int(p_lo), C.c_int(p_hi),
                           _p(stack))
        for p in range(p_lo, p_hi):
            F = scipy.fft.ifft2(stack[p - p_lo], norm="forward", workers=nthreads, overwrite_x=True)
            L.oracle_fast_wscreen(C.byref(cp), _p(np.ascontiguousarray(F, dtype=np.complex64)), C.c_int(p), _p(img))
    cu = kernel_ft(pl, (np.arange(nx) - nx // 2) / pl.nu)
    cv = kernel_ft(pl, (np.arange(ny) - ny // 2) / pl.nv)
    ntab = 4097
    tab_x = np.linspace(pl.nm1_min, 0.0, ntab)
    cw = kernel_ft(pl, (tab_x + pl.nshift) * pl.dw)
    L.oracle_fast_correct(C.byref(cp), _p(cu), _p(cv), _p(cw), C.c_int(ntab), C.c_double(pl.nm1_min), _p(img))
    return img
