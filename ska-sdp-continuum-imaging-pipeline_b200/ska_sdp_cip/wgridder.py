"""B200 replacement of ``ducc0.wgridder.ms2dirty`` -- the one native call of the reference
(src/ska_sdp_cip/invert.py:13, :170-183).

Same positional arguments and keywords as the ducc0 function the reference calls, so
``ducc_invert`` below reads like the reference's.  Inputs may be numpy arrays (host memory;
copied to the GPU inside the call) or CUDA torch tensors (used in place, result returned as a
CUDA tensor).  All arithmetic happens in libcip_b200.so; nothing here computes.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _cabi
from ._cabi import Context, InvertArgs, Plan, Report, ScanResult, check, lib


def _is_torch_cuda(a) -> bool:
    return a is not None and not isinstance(a, np.ndarray) and hasattr(a, "is_cuda") and a.is_cuda


def _is_c128(a) -> bool:
    return a is not None and str(a.dtype) in ("complex128", "torch.complex128")


def _prep_host(uvw, freq, ms, wgt, mask):
    uvw = np.ascontiguousarray(uvw, dtype=np.float64)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    if uvw.ndim != 2 or uvw.shape[1] != 3:
        raise ValueError(f"uvw must have shape (nrow, 3), got {uvw.shape}")
    ms = np.asarray(ms)
    if ms.ndim != 2 or ms.shape != (uvw.shape[0], freq.shape[0]):
        raise ValueError(f"ms must have shape (nrow, nchan) = {(uvw.shape[0], freq.shape[0])}, got {ms.shape}")
    # complex128 visibilities go through as they are, with float64 weights (ducc0's double-precision call)
    fp64_in = ms.dtype == np.complex128
    ms = np.ascontiguousarray(ms, dtype=np.complex128 if fp64_in else np.complex64)
    if wgt is not None:
        wgt = np.ascontiguousarray(wgt, dtype=np.float64 if fp64_in else np.float32)
        if wgt.shape != ms.shape:
            raise ValueError("wgt must have the shape of ms")
    if mask is not None:
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        if mask.shape != ms.shape:
            raise ValueError("mask must have the shape of ms")
    return uvw, freq, ms, wgt, mask, fp64_in


def _host_result(shape, dtype):
    """A new array for the dirty image.  Large images come from torch's caching pinned-memory
    allocator (a numpy view that keeps the block alive and returns it to the cache when collected):
    the device -> host copy then runs at full PCIe speed and a fresh result does not page-fault."""
    if int(np.prod(shape)) * np.dtype(dtype).itemsize >= (1 << 20):
        try:
            import torch
            return torch.empty(shape, dtype=torch.float64 if dtype == np.float64 else torch.float32,
                               pin_memory=True).numpy()
        except Exception:  # noqa: BLE001  (no torch / no pinned memory: plain pageable memory works too)
            pass
    return np.empty(shape, dtype=dtype)


def make_args(uvw, freq, ms, wgt, mask, npix_x, npix_y, pixsize_x, pixsize_y, epsilon, do_wstacking, fp64,
              memspace, w_slab=None) -> InvertArgs:
    a = InvertArgs()
    a.uvw = _cabi._ptr(uvw)
    a.nrow = int(uvw.shape[0])
    a.freq = _cabi._ptr(freq)
    a.nchan = int(freq.shape[0])
    a.vis = _cabi._ptr(ms)
    a.wgt = _cabi._ptr(wgt)
    a.mask = _cabi._ptr(mask)
    a.vis_row_stride = 0
    a.wgt_row_stride = 0
    a.npix_x, a.npix_y = int(npix_x), int(npix_y)
    a.pixsize_x, a.pixsize_y = float(pixsize_x), float(pixsize_y)
    a.epsilon = float(epsilon)
    a.do_wstacking = int(bool(do_wstacking))
    a.accumulate_fp64 = int(bool(fp64))
    a.memspace = memspace
    a.w_lo, a.w_hi = (float(w_slab[0]), float(w_slab[1])) if w_slab else (0.0, 0.0)
    a.vis_c128 = int(_is_c128(ms))
    return a


def make_jagged_args(uvw, freq, vis_flat, row_offsets, chan_start, chan_stop, wgt_flat, npix_x, npix_y, pixsize_x,
                     pixsize_y, epsilon, do_wstacking, fp64, memspace) -> InvertArgs:
    """Arguments for rows that hold only a range of channels each (the uvw tile layout): row r = channels
    [chan_start[r], chan_stop[r]) at elements row_offsets[r] ... of the flat vis (and wgt)."""
    a = make_args(uvw, freq, vis_flat, wgt_flat, None, npix_x, npix_y, pixsize_x, pixsize_y, epsilon, do_wstacking, fp64,
                  memspace)
    a.row_vis_offset = _cabi._ptr(row_offsets)
    a.row_chan_start = _cabi._ptr(chan_start)
    a.row_chan_stop = _cabi._ptr(chan_stop)
    a.nvis = int(vis_flat.shape[0])
    return a


def tiles2dirty(tiles, freq, npix_x, npix_y, pixsize_x, pixsize_y, epsilon=1e-4, do_wstacking=True,
                double_precision_accumulation=False, *, device: int = -1, return_info: bool = False):
    """Dirty image of the visibilities held in uvw tiles (``ska_sdp_cip.uvw_tiling.Tile``: uvw per row slice, flat
    visibilities, channel_start_indices / channel_stop_indices), gridded as they are stored -- the link between the
    reorder step and the gridder that the reference has not made (src/ska_sdp_cip/uvw_tiling/reorder.py:129-155
    writes the tiles, nothing reads them).  Tiles carry no weights: every sample has weight 1.  ``freq`` are the
    channel frequencies the tiles' channel indices refer to."""
    tiles = list(tiles)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    cat = lambda name, dt: (np.ascontiguousarray(np.concatenate([np.asarray(getattr(t, name)) for t in tiles]), dtype=dt)
                            if tiles else np.zeros(0, dtype=dt))
    uvw = cat("uvw", np.float64).reshape(-1, 3)
    c0, c1 = cat("channel_start_indices", np.int32), cat("channel_stop_indices", np.int32)
    vis = cat("visibilities", np.complex64)
    if len(c0) and (c0.min() < 0 or c1.max() > len(freq) or np.any(c1 < c0)):
        raise ValueError("tile channel ranges do not fit the given channel frequencies")
    lens = (c1 - c0).astype(np.int64)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64) if len(lens) else np.zeros(0, dtype=np.int64)
    if int(lens.sum()) != len(vis):
        raise ValueError("tile visibilities do not match their channel ranges")
    fp64 = bool(double_precision_accumulation)
    out = _host_result((npix_x, npix_y), np.float64 if fp64 else np.float32)
    args = make_jagged_args(uvw, freq, vis, off, c0, c1, None, npix_x, npix_y, pixsize_x, pixsize_y, epsilon,
                            do_wstacking, fp64, 0)
    ctx = Context.get(device)
    scan, plan, rep = ScanResult(), Plan(), Report()
    with ctx.lock:
        check(lib().cip_invert(ctx.handle, C.byref(args), _cabi._ptr(out), C.byref(scan), C.byref(plan), C.byref(rep)))
    if return_info:
        return out, {"scan": {n: getattr(scan, n) for n, _ in scan._fields_}, "plan": plan.as_dict(),
                     "report": rep.as_dict()}
    return out


def ms2dirty(uvw, freq, ms, wgt=None, npix_x=0, npix_y=0, pixsize_x=0.0, pixsize_y=0.0, nu=0, nv=0,
             epsilon=1e-4, do_wstacking=False, nthreads=1, verbosity=0, mask=None,
             double_precision_accumulation=False, *, device: int = -1, return_info: bool = False, w_slab=None):
    """Dirty image (npix_x, npix_y) of the given visibilities; see the module docstring.

    ``nthreads``, ``nu``, ``nv`` and ``verbosity`` are accepted for signature compatibility and ignored.
    complex128 ``ms`` or ``double_precision_accumulation=True`` selects float64 grids and a float64 result.
    ``w_slab=(w_lo, w_hi)`` (an extension) restricts the image to the samples with w_lo <= |w| < w_hi wavelengths:
    one share of a |w|-sharded invert (SURVEY.md 8e option 3).
    """
    del nthreads, nu, nv, verbosity
    dev_in = _is_torch_cuda(ms)
    if dev_in:
        import torch
        for t in (uvw, freq, ms):
            if not (_is_torch_cuda(t) and t.is_contiguous()):
                raise ValueError("with CUDA inputs, uvw, freq and ms must all be contiguous CUDA tensors")
        if uvw.dtype != torch.float64 or freq.dtype != torch.float64 or ms.dtype not in (torch.complex64, torch.complex128):
            raise ValueError("CUDA inputs must be float64 uvw/freq and complex64 or complex128 ms")
        wdt = torch.float64 if ms.dtype == torch.complex128 else torch.float32
        if wgt is not None and (not _is_torch_cuda(wgt) or wgt.dtype != wdt or not wgt.is_contiguous()):
            raise ValueError(f"wgt must be a contiguous {wdt} CUDA tensor (the real type of ms)")
        if mask is not None and (not _is_torch_cuda(mask) or mask.dtype != torch.uint8 or not mask.is_contiguous()):
            raise ValueError("mask must be a contiguous uint8 CUDA tensor")
        fp64 = bool(double_precision_accumulation) or ms.dtype == torch.complex128
        device = ms.device.index
        out = torch.empty((npix_x, npix_y), dtype=torch.float64 if fp64 else torch.float32, device=ms.device)
        torch.cuda.current_stream(ms.device).synchronize()
        memspace = 1
    else:
        uvw, freq, ms, wgt, mask, fp64_in = _prep_host(uvw, freq, ms, wgt, mask)
        fp64 = bool(double_precision_accumulation) or fp64_in
        out = _host_result((npix_x, npix_y), np.float64 if fp64 else np.float32)
        memspace = 0
    args = make_args(uvw, freq, ms, wgt, mask, npix_x, npix_y, pixsize_x, pixsize_y, epsilon, do_wstacking, fp64,
                     memspace, w_slab=w_slab)
    ctx = Context.get(device)
    scan, plan, rep = ScanResult(), Plan(), Report()
    with ctx.lock:
        check(lib().cip_invert(ctx.handle, C.byref(args), _cabi._ptr(out), C.byref(scan), C.byref(plan), C.byref(rep)))
    if return_info:
        return out, {"scan": {n: getattr(scan, n) for n, _ in scan._fields_}, "plan": plan.as_dict(),
                     "report": rep.as_dict()}
    return out
