"""ctypes binding of libcip_b200.so (the C-ABI declared in include/cip.h).

This is the only place the package crosses into native code.  There is no CPU
fallback: if the shared library is missing or no CUDA device is present the
calls raise.  PyTorch tensors (or numpy arrays) are used purely as buffers;
only raw pointers and sizes cross the boundary.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path

import numpy as np

_LIB_PATH = Path(__file__).resolve().parent / "_lib" / "libcip_b200.so"
_lib = None
_lock = threading.Lock()


class CipError(RuntimeError):
    """Raised for CUDA / cuFFT / memory failures (ducc0 raises RuntimeError too)."""


class InvertArgs(C.Structure):
    _fields_ = [("uvw", C.c_void_p), ("nrow", C.c_int64), ("freq", C.c_void_p), ("nchan", C.c_int64),
                ("vis", C.c_void_p), ("wgt", C.c_void_p), ("mask", C.c_void_p),
                ("vis_row_stride", C.c_int64), ("wgt_row_stride", C.c_int64),
                ("npix_x", C.c_int), ("npix_y", C.c_int), ("pixsize_x", C.c_double), ("pixsize_y", C.c_double),
                ("epsilon", C.c_double), ("do_wstacking", C.c_int), ("accumulate_fp64", C.c_int),
                ("memspace", C.c_int), ("row_vis_offset", C.c_void_p), ("row_chan_start", C.c_void_p),
                ("row_chan_stop", C.c_void_p), ("nvis", C.c_int64), ("w_lo", C.c_double), ("w_hi", C.c_double),
                ("vis_c128", C.c_int32)]


class ScanResult(C.Structure):
    _fields_ = [("wmin", C.c_double), ("wmax", C.c_double), ("nactive", C.c_int64), ("sum_wgt", C.c_double)]


class Plan(C.Structure):
    _fields_ = [("npix_x", C.c_int), ("npix_y", C.c_int), ("pixsize_x", C.c_double), ("pixsize_y", C.c_double),
                ("sigma", C.c_double), ("support", C.c_int), ("beta", C.c_double), ("nu", C.c_int), ("nv", C.c_int),
                ("nshift", C.c_double), ("dw", C.c_double), ("w0", C.c_double), ("nplanes", C.c_int),
                ("accumulate_fp64", C.c_int), ("do_wstacking", C.c_int), ("plane_block", C.c_int)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Report(C.Structure):
    _fields_ = [("ms_total", C.c_double), ("ms_h2d", C.c_double), ("ms_scan", C.c_double), ("ms_bin", C.c_double),
                ("ms_grid", C.c_double), ("ms_fft", C.c_double), ("ms_wscreen", C.c_double),
                ("ms_correct", C.c_double), ("ms_d2h", C.c_double), ("launches", C.c_int64),
                ("nactive", C.c_int64), ("nblocks", C.c_int), ("ms_zero", C.c_double)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def lib_path() -> Path:
    return Path(os.environ.get("CIP_B200_LIB", _LIB_PATH))


def lib():
    """Load libcip_b200.so; raises ImportError with build instructions if it is missing."""
    global _lib
    with _lock:
        if _lib is None:
            path = lib_path()
            if not path.exists():
                raise ImportError(
                    f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(or `make -C ska-sdp-continuum-imaging-pipeline_b200`); there is no CPU fallback")
            L = C.CDLL(str(path))
            L.cip_last_error.restype = C.c_char_p
            L.cip_version.restype = C.c_char_p
            L.cip_ctx_stream.restype = C.c_void_p
            L.cip_ctx_stream.argtypes = [C.c_void_p]
            L.cip_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
            L.cip_ctx_destroy.argtypes = [C.c_void_p]
            L.cip_ctx_synchronize.argtypes = [C.c_void_p]
            L.cip_ctx_launches.argtypes = [C.c_void_p]
            L.cip_ctx_launches.restype = C.c_int64
            L.cip_ctx_device.argtypes = [C.c_void_p]
            L.cip_ctx_device.restype = C.c_int
            L.cip_current_device.argtypes = [C.POINTER(C.c_int)]
            L.cip_plan_make.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int,
                                        C.c_double, C.c_double, C.c_int64, C.c_double, C.c_int, C.POINTER(Plan)]
            L.cip_kernel_table.argtypes = [C.c_double, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
            L.cip_kernel_ft.argtypes = [C.POINTER(Plan), C.c_void_p, C.c_int64, C.c_void_p]
            L.cip_scan.argtypes = [C.c_void_p, C.POINTER(InvertArgs), C.POINTER(ScanResult)]
            L.cip_grid_accumulate.argtypes = [C.c_void_p, C.POINTER(InvertArgs), C.POINTER(Plan), C.c_void_p,
                                              C.POINTER(Report)]
            L.cip_w_histogram.argtypes = [C.c_void_p, C.POINTER(InvertArgs), C.c_double, C.c_double, C.c_int, C.c_void_p]
            L.cip_w_histogram_static.argtypes = L.cip_w_histogram.argtypes
            L.cip_grid_accumulate_many.argtypes = [C.c_void_p, C.POINTER(InvertArgs), C.c_int, C.POINTER(Plan), C.c_void_p,
                                                   C.POINTER(Report)]
            L.cip_bin_inspect.argtypes = [C.c_void_p, C.POINTER(InvertArgs), C.POINTER(Plan), C.POINTER(C.c_int64),
                                          C.POINTER(C.c_int64), C.POINTER(C.c_int), C.c_void_p, C.c_void_p]
            L.cip_plane_fft_crop.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
            L.cip_correct_image.argtypes = [C.c_void_p, C.POINTER(Plan), C.c_void_p, C.c_int, C.c_int, C.c_double]
            L.cip_invert.argtypes = [C.c_void_p, C.POINTER(InvertArgs), C.c_void_p, C.POINTER(ScanResult),
                                     C.POINTER(Plan), C.POINTER(Report)]
            L.cip_stokes_i.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                       C.c_void_p, C.c_void_p, C.c_int]
            L.cip_stokes_i_async.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                             C.c_void_p, C.c_void_p]
            L.cip_tile_bin.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                       C.POINTER(C.c_double * 3), C.c_int64, C.c_int, C.POINTER(C.c_void_p),
                                       C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
            L.cip_tile_fetch.argtypes = [C.c_void_p] * 7
            L.cip_tile_gather.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                          C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
            L.cip_tile_free.argtypes = [C.c_void_p, C.c_void_p]
            _lib = L
    return _lib


EXPORTS = ["cip_last_error", "cip_version", "cip_ctx_create", "cip_ctx_destroy", "cip_ctx_stream",
           "cip_ctx_synchronize", "cip_ctx_launches", "cip_ctx_device", "cip_current_device", "cip_plan_make", "cip_kernel_table", "cip_kernel_ft", "cip_scan", "cip_w_histogram", "cip_w_histogram_static",
           "cip_grid_accumulate", "cip_grid_accumulate_many", "cip_bin_inspect", "cip_plane_fft_crop", "cip_correct_image", "cip_invert", "cip_stokes_i", "cip_stokes_i_async", "cip_tile_bin",
           "cip_tile_fetch", "cip_tile_gather", "cip_tile_free"]


def check(status: int) -> None:
    """CIP_EINVAL -> ValueError, everything else -> CipError (a RuntimeError), as SURVEY 8(b) specifies."""
    if status == 0:
        return
    msg = lib().cip_last_error().decode(errors="replace")
    if status == 1:
        raise ValueError(msg)
    if status == 2:
        raise MemoryError(msg)
    raise CipError(f"libcip_b200 status {status}: {msg}")


def current_device() -> int:
    """Index of the calling thread's current CUDA device (what ``device=-1`` means)."""
    dev = C.c_int(-1)
    check(lib().cip_current_device(C.byref(dev)))
    return dev.value


class Context:
    """One context per process and GPU (cip_ctx is not re-entrant; calls are serialised by a lock)."""

    _instances: dict[int, "Context"] = {}
    _create_lock = threading.Lock()

    def __init__(self, device: int = -1):
        self._h = C.c_void_p()
        check(lib().cip_ctx_create(device, C.byref(self._h)))
        self.device = int(lib().cip_ctx_device(self._h))
        self.lock = threading.Lock()

    @classmethod
    def get(cls, device: int = -1) -> "Context":
        """The context of ``device`` (-1 or None: the current CUDA device, resolved to its index first, so that
        a process never holds two contexts -- two plane stacks -- on one GPU)."""
        if device is None or device < 0:
            device = current_device()
        with cls._create_lock:
            inst = cls._instances.get(device)
            if inst is None:
                inst = cls._instances[device] = cls(device)
        return inst

    @property
    def handle(self):
        return self._h

    def close(self):
        if self._h:
            lib().cip_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


def _ptr(a):
    """Raw pointer of a numpy array or torch tensor (None -> NULL)."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


def plan_make(nx, ny, px, py, epsilon, wmin, wmax, nactive, *, do_wstacking=True, fp64=False, sigma=0.0,
              support=0) -> Plan:
    p = Plan()
    check(lib().cip_plan_make(nx, ny, px, py, epsilon, int(do_wstacking), int(fp64), wmin, wmax, int(nactive),
                              float(sigma or 0.0), int(support or 0), C.byref(p)))
    return p


JOB_DTYPE = np.dtype([("group_begin", "<u4"), ("ngroups", "<u4"), ("tu", "<u2"), ("tv", "<u2"), ("win", "<u2"),
                      ("lu", "u1"), ("lv", "u1")])


def bin_inspect(ctx: "Context", args: InvertArgs, plan: Plan):
    """(jobs structured array, records as bytes (ngroups * 8, rec_bytes)) of stage G1 -- see cip_bin_inspect."""
    nj, ng, rb = C.c_int64(), C.c_int64(), C.c_int()
    with ctx.lock:
        check(lib().cip_bin_inspect(ctx.handle, C.byref(args), C.byref(plan), C.byref(nj), C.byref(ng), C.byref(rb), None,
                                    None))
        jobs = np.zeros(nj.value, dtype=JOB_DTYPE)
        recs = np.zeros((ng.value * 8, rb.value), dtype=np.uint8)
        check(lib().cip_bin_inspect(ctx.handle, C.byref(args), C.byref(plan), C.byref(nj), C.byref(ng), C.byref(rb),
                                    jobs.ctypes.data, recs.ctypes.data))
    return jobs, recs


def plane_fft_crop(ctx: "Context", plane, nx: int, ny: int, overwrite: bool = False):
    """Retained nx x ny cells of the unnormalised e^{+i} 2-D transform of a CUDA complex64 / complex128 plane
    (torch tensor, shape (nu, nv)), in image order -- see cip_plane_fft_crop.  The library works in place on the
    plane; unless ``overwrite`` a copy is transformed."""
    import torch
    assert plane.is_cuda and plane.is_contiguous() and plane.dtype in (torch.complex64, torch.complex128)
    nu, nv = plane.shape
    if not overwrite:
        plane = plane.clone()
    out = torch.empty((nx, ny), dtype=plane.dtype, device=plane.device)
    torch.cuda.current_stream(plane.device).synchronize()
    with ctx.lock:
        check(lib().cip_plane_fft_crop(ctx.handle, plane.data_ptr(), nu, nv, nx, ny, int(plane.dtype == torch.complex128),
                                       out.data_ptr()))
    return out


def kernel_ft(plan: Plan, xi) -> np.ndarray:
    xi = np.ascontiguousarray(xi, dtype=np.float64)
    out = np.empty_like(xi)
    check(lib().cip_kernel_ft(C.byref(plan), xi.ctypes.data, xi.size, out.ctypes.data))
    return out
