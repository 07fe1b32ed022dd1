"""Drop-in for ``ska_sdp_cip.invert`` (reference: src/ska_sdp_cip/invert.py) with the native step
on a B200: same functions, same arguments, same return values.

Host Python keeps what the reference keeps on the host (reading chunks, fan-out over chunks);
everything numerical crosses the C-ABI of include/cip.h:
  Stokes-I / effective weights (invert.py:72-116)   -> cip_stokes_i
  ducc0.wgridder.ms2dirty      (invert.py:170-183)   -> cip_invert, or cip_scan + cip_plan_make +
                                                        cip_grid_accumulate + cip_correct_image
  effective_weights.sum()      (invert.py:184)       -> cip_scan_result.sum_wgt
  integrate_weighted_images    (invert.py:200-209)   -> images are summed uncorrected on the GPU
                                                        (NCCL reduce across ranks), then corrected once
PyTorch supplies device buffers and the process group only.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Any, Iterable, Optional, Sequence

import numpy as np
from numpy.typing import NDArray

from . import _cabi
from ._cabi import Context, Plan, Report, ScanResult, check, lib
from .measurement_set import MeasurementSetReader
from .wgridder import make_args, ms2dirty

EPSILON = 1e-4  # hard-coded in the reference (invert.py:179)


def _torch():
    import torch  # deferred: importing torch takes a while and the tiling API does not need it
    return torch


def _device(device: Optional[int] = None):
    torch = _torch()
    if not torch.cuda.is_available():
        raise _cabi.CipError("no CUDA device available; this package has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device() if device is None else device)


@dataclass
class StokesIGridderInput:
    """Stokes I visibilities and associated arrays handed to the gridder (reference: invert.py:40-116).

    Built by ``from_measurement_set_reader`` the arrays are CUDA tensors and ``weights`` already holds
    the effective weights (zero where flagged), so ``flags`` is None; built by hand from numpy arrays
    it behaves exactly like the reference's dataclass.
    """

    channel_frequencies: Any
    flags: Any
    uvw: Any
    visibilities: Any
    weights: Any

    def effective_weights(self):
        """``weights x (1 - flags)`` (invert.py:72-76)."""
        if self.flags is None:
            return self.weights
        if isinstance(self.weights, np.ndarray):
            return np.logical_not(self.flags) * self.weights
        return self.flags.logical_not() * self.weights

    @classmethod
    def from_measurement_set_reader(cls, ms_reader: MeasurementSetReader, *, device: Optional[int] = None
                                    ) -> "StokesIGridderInput":
        """Read one chunk and convert to Stokes I on the GPU: vis_I = (v0+v3)/2,
        w_I = 4/(1/w0+1/w3), flagged if either hand is flagged (invert.py:78-116)."""
        torch = _torch()
        dev = _device(device)
        vis = np.ascontiguousarray(ms_reader.visibilities(), dtype=np.complex64)
        flags = np.ascontiguousarray(ms_reader.flags(), dtype=np.bool_)
        weights = np.ascontiguousarray(ms_reader.weights(), dtype=np.float32)
        if vis.ndim != 3 or vis.shape[2] != 4:
            raise ValueError("expected 4 polarisation products")
        nrow, nchan = vis.shape[:2]
        d_vis4 = torch.from_numpy(vis).to(dev, non_blocking=False)
        d_flags4 = torch.from_numpy(flags.view(np.uint8)).to(dev)
        d_wgt4 = torch.from_numpy(weights).to(dev)
        d_vis = torch.empty((nrow, nchan), dtype=torch.complex64, device=dev)
        d_effw = torch.empty((nrow, nchan), dtype=torch.float32, device=dev)
        torch.cuda.current_stream(dev).synchronize()
        ctx = Context.get(dev.index)
        with ctx.lock:
            check(lib().cip_stokes_i(ctx.handle, d_vis4.data_ptr(), d_flags4.data_ptr(), d_wgt4.data_ptr(),
                                     nrow, nchan, d_vis.data_ptr(), d_effw.data_ptr(), 1))
        freqs = torch.from_numpy(np.ascontiguousarray(ms_reader.channel_frequencies(), dtype=np.float64)).to(dev)
        uvw = torch.from_numpy(np.ascontiguousarray(ms_reader.uvw(), dtype=np.float64)).to(dev)
        return cls(freqs, None, uvw, d_vis, d_effw)


def _pixel_size_lm(pixel_size_asec: float) -> float:
    return float(np.sin(np.radians(pixel_size_asec / 3600.0)))  # invert.py:163


def invert_measurement_set(ms_reader: MeasurementSetReader, num_pixels: int, pixel_size_asec: float,
                           nthreads: int = os.cpu_count(), *, epsilon: float = EPSILON,
                           accumulate_fp64: bool = False) -> NDArray:
    """Invert the given measurement set, returning a dirty image (reference: invert.py:119-149).

    ``nthreads`` is accepted for compatibility and ignored (the work runs on the GPU).
    """
    gridding_input = StokesIGridderInput.from_measurement_set_reader(ms_reader)
    image, total_weight = ducc_invert(gridding_input, num_pixels, pixel_size_asec, nthreads=nthreads,
                                      epsilon=epsilon, accumulate_fp64=accumulate_fp64)
    return (1.0 / total_weight) * image


def ducc_invert(gridder_input: StokesIGridderInput, num_pixels: int, pixel_size_asec: float,
                nthreads: int = os.cpu_count(), *, epsilon: float = EPSILON,
                accumulate_fp64: bool = False) -> tuple[NDArray, float]:
    """Unscaled dirty image and its total gridding weight (reference: invert.py:152-184).
    The name is kept for drop-in compatibility; the gridder is libcip_b200, not ducc0."""
    pixel_size_lm = _pixel_size_lm(pixel_size_asec)
    effective_weights = gridder_input.effective_weights()
    image, info = ms2dirty(gridder_input.uvw, gridder_input.channel_frequencies, gridder_input.visibilities,
                           effective_weights, num_pixels, num_pixels, pixel_size_lm, pixel_size_lm,
                           epsilon=epsilon, do_wstacking=True, nthreads=nthreads, mask=None,
                           double_precision_accumulation=accumulate_fp64, return_info=True)
    total = np.float32(info["scan"]["sum_wgt"])
    if not isinstance(image, np.ndarray):
        image = image.cpu().numpy()
    return image, total


def worker_ducc_invert(gridder_input: StokesIGridderInput, num_pixels: int, pixel_size_asec: float, *,
                       epsilon: float = EPSILON, accumulate_fp64: bool = False) -> tuple[NDArray, float]:
    """Same as ducc_invert, for use as a dask task (reference: invert.py:187-197)."""
    return ducc_invert(gridder_input, num_pixels, pixel_size_asec, epsilon=epsilon, accumulate_fp64=accumulate_fp64)


def integrate_weighted_images(weighted_images: Iterable[tuple[NDArray, float]]) -> NDArray:
    """Sum of chunk images over sum of chunk weights (reference: invert.py:200-209)."""
    images, weights = zip(*weighted_images)
    return sum(images) / sum(weights)


def shard_of_rank(chunks: Sequence, rank: int, world: int) -> list:
    """Chunks of the reference's partition (row-major, frequency-minor) owned by ``rank``: round-robin,
    so that with freq_chunks = world each rank gets one frequency chunk of every row chunk."""
    return list(chunks[rank::world])


def agree_on_shard_statistics(wmin: float, wmax: float, nactive: int, sumw: float, *, group=None, device=None):
    """All ranks must build their partial images with the same (sigma, W, dw, nshift, nplanes): combine
    the per-rank |w| range (min/max), active count and weight sum (sums) over the process group.
    Two small all-reduces; a no-op without an initialised process group."""
    if not _dist_ready():
        return wmin, wmax, nactive, sumw
    torch = _torch()
    dist = torch.distributed
    if dist.get_world_size(group) == 1:
        return wmin, wmax, nactive, sumw
    t = torch.tensor([-wmin, wmax], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    s = torch.tensor([float(nactive), sumw], dtype=torch.float64, device=device)
    dist.all_reduce(s, op=dist.ReduceOp.SUM, group=group)
    return -float(t[0]), float(t[1]), int(s[0].item()), float(s[1])


def reduce_image_rows(image, *, group=None):
    """Sum the partial images over the ranks, leaving each rank with a contiguous block of rows:
    returns (rows tensor, row_begin, row_end).  NCCL: one reduce-scatter.  Images whose row count is
    not divisible by the world size (and the gloo backend, which has no reduce-scatter) use an
    all-reduce and every rank keeps all rows."""
    torch = _torch()
    n = image.shape[0]
    if not _dist_ready() or torch.distributed.get_world_size(group) == 1:
        return image, 0, n
    dist = torch.distributed
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if n % world == 0 and dist.get_backend(group) == "nccl":
        rows = n // world
        mine = torch.empty((rows, image.shape[1]), dtype=image.dtype, device=image.device)
        dist.reduce_scatter_tensor(mine, image, op=dist.ReduceOp.SUM, group=group)
        return mine, rank * rows, (rank + 1) * rows
    dist.all_reduce(image, op=dist.ReduceOp.SUM, group=group)
    if n % world == 0:
        rows = n // world
        return image[rank * rows:(rank + 1) * rows], rank * rows, (rank + 1) * rows
    return image, 0, n


def gather_image_rows(image, mine, row_begin: int, row_end: int, *, group=None):
    """Inverse of reduce_image_rows after the per-block correction: the full image on every rank."""
    if row_begin == 0 and row_end == image.shape[0]:
        return mine
    _torch().distributed.all_gather_into_tensor(image, mine.contiguous(), group=group)
    return image


# ---- sharded invert: the B200 form of dask_invert_measurement_set ------------------------------

def accumulate_shards(shards: Sequence[StokesIGridderInput], num_pixels: int, pixel_size_lm: float, *,
                      epsilon: float = EPSILON, accumulate_fp64: bool = False, group=None,
                      device: Optional[int] = None, reports: Optional[list] = None):
    """Grid this rank's shards into ONE uncorrected partial image on the GPU.

    All ranks of ``group`` (a torch.distributed process group, or None for a single process) first
    agree on the |w| range and the active count so that every partial image is built with the same
    (sigma, W, dw, nshift); partial images are then linear in the visibilities and can be summed
    before the correction.  Returns (image CUDA tensor, Plan, sum of weights).
    """
    torch = _torch()
    dev = _device(device)
    ctx = Context.get(dev.index)
    L = lib()
    args_list, wmin, wmax, nactive, sumw = [], np.inf, 0.0, 0, 0.0
    torch.cuda.current_stream(dev).synchronize()
    for sh in shards:
        if isinstance(sh.visibilities, np.ndarray):
            # host shard: handed to the library as host memory -- cip_scan stages weights and uvw only,
            # cip_grid_accumulate reuses them and streams the visibilities in behind the binning/gridding
            c = lambda a, dt: np.ascontiguousarray(a, dtype=dt)
            sh = StokesIGridderInput(c(sh.channel_frequencies, np.float64), None, c(sh.uvw, np.float64),
                                     c(sh.visibilities, np.complex64), c(sh.effective_weights(), np.float32))
        effw = sh.effective_weights()
        a = make_args(sh.uvw, sh.channel_frequencies, sh.visibilities, effw, None, num_pixels, num_pixels,
                      pixel_size_lm, pixel_size_lm, epsilon, True, accumulate_fp64,
                      0 if isinstance(sh.visibilities, np.ndarray) else 1)
        scan = ScanResult()
        with ctx.lock:
            check(L.cip_scan(ctx.handle, C.byref(a), C.byref(scan)))
        args_list.append((a, sh, effw))  # keep the buffers alive
        if scan.nactive:
            wmin, wmax = min(wmin, scan.wmin), max(wmax, scan.wmax)
        nactive += scan.nactive
        sumw += scan.sum_wgt
    wmin, wmax, nactive, sumw = agree_on_shard_statistics(wmin, wmax, nactive, sumw, group=group, device=dev)
    if nactive == 0:
        wmin = wmax = 0.0
    plan = _cabi.plan_make(num_pixels, num_pixels, pixel_size_lm, pixel_size_lm, epsilon, wmin, wmax, nactive,
                           do_wstacking=True, fp64=accumulate_fp64,
                           sigma=float(os.environ.get("CIP_SIGMA", 0) or 0),
                           support=int(os.environ.get("CIP_SUPPORT", 0) or 0))
    image = torch.zeros((num_pixels, num_pixels), dtype=torch.float64 if accumulate_fp64 else torch.float32,
                        device=dev)
    torch.cuda.current_stream(dev).synchronize()
    if nactive:
        for a, _sh, _w in args_list:
            rep = Report()
            with ctx.lock:
                check(L.cip_grid_accumulate(ctx.handle, C.byref(a), C.byref(plan), image.data_ptr(), C.byref(rep)))
            if reports is not None:
                reports.append(rep.as_dict())
    return image, plan, sumw


def _dist_ready() -> bool:
    torch = _torch()
    return torch.distributed.is_available() and torch.distributed.is_initialized()


def reduce_and_correct(image, plan: Plan, total_weight: float, *, group=None, normalise: bool = True):
    """Sum the partial images of all ranks and apply the grid correction once (P5 + G4 final).

    Multi-rank: reduce-scatter by image rows over NCCL, each rank corrects its row block
    (cip_correct_image with the 1/sum(weights) scale folded in), all-gather.  Returns the full
    corrected image on every rank."""
    torch = _torch()
    ctx = Context.get(image.device.index)
    scale = (1.0 / total_weight) if normalise else 1.0
    torch.cuda.current_stream(image.device).synchronize()
    mine, r0, r1 = reduce_image_rows(image, group=group)
    torch.cuda.current_stream(image.device).synchronize()
    # cip_correct_image takes the pointer to row 0 of the full image: offset back by this rank's first row
    base = mine.data_ptr() - r0 * image.shape[1] * image.element_size()
    with ctx.lock:
        check(lib().cip_correct_image(ctx.handle, C.byref(plan), base, r0, r1, scale))
    return gather_image_rows(image, mine, r0, r1, group=group)


def dask_invert_measurement_set(ms_reader: MeasurementSetReader, client, num_pixels: int, pixel_size_asec: float,
                                *, row_chunks: Optional[int] = 1, freq_chunks: Optional[int] = None,
                                epsilon: float = EPSILON, accumulate_fp64: bool = False) -> NDArray:
    """Invert the measurement set chunk by chunk (reference: invert.py:212-270).

    The partition is the reference's: ``ms_reader.partition(row_chunks, freq_chunks)``, row-major and
    frequency-minor.  With a dask ``client`` the chunks are read and inverted as dask tasks exactly like
    the reference (each worker needs a GPU).  With ``client=None`` the chunks are sharded round-robin
    over the ranks of the initialised torch.distributed process group (one process per GPU; a single
    process takes them all), gridded into one partial image per GPU and combined by a single NCCL
    reduction before the correction.
    """
    row_chunks = max(row_chunks or 1, 1)
    if not freq_chunks:
        if client is not None:
            num_workers = len(client.scheduler_info()["workers"])
        else:
            num_workers = _torch().distributed.get_world_size() if _dist_ready() else 1
        freq_chunks = min(ms_reader.num_channels, max(num_workers, 1))
    chunks = ms_reader.partition(row_chunks, freq_chunks)

    if client is not None:
        weighted_images = []
        for chunk in chunks:
            gridder_input = client.submit(StokesIGridderInput.from_measurement_set_reader, chunk)
            weighted_images.append(client.submit(worker_ducc_invert, gridder_input, num_pixels, pixel_size_asec,
                                                 epsilon=epsilon, accumulate_fp64=accumulate_fp64,
                                                 resources={"processing_slots": 1}))
        return client.submit(integrate_weighted_images, weighted_images).result()

    torch = _torch()
    rank, world = (torch.distributed.get_rank(), torch.distributed.get_world_size()) if _dist_ready() else (0, 1)
    shards = [StokesIGridderInput.from_measurement_set_reader(c) for c in shard_of_rank(chunks, rank, world)]
    image, plan, sumw = accumulate_shards(shards, num_pixels, _pixel_size_lm(pixel_size_asec), epsilon=epsilon,
                                          accumulate_fp64=accumulate_fp64)
    image = reduce_and_correct(image, plan, float(np.float32(sumw)))
    return image.cpu().numpy()
