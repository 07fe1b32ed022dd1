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
    def from_measurement_set_reader(cls, ms_reader: MeasurementSetReader, *, device: Optional[int] = None,
                                    upload_rows: Optional[int] = None) -> "StokesIGridderInput":
        """Read one chunk and convert to Stokes I on the GPU: vis_I = (v0+v3)/2,
        w_I = 4/(1/w0+1/w3), flagged if either hand is flagged (invert.py:78-116).

        The chunk is read and uploaded in row blocks of ``upload_rows`` (default: about 256 MB of 4-polarisation
        data) through two pinned staging buffers: while block k is converted on the GPU (cip_stokes_i) block k+1 is
        read from the reader, copied into the other pinned buffer and sent host -> device on a copy stream.  Only
        the Stokes-I visibilities and effective weights (12 of the 52 bytes per sample) stay resident."""
        torch = _torch()
        dev = _device(device)
        nrow_total, nchan = int(ms_reader.num_data_rows), int(ms_reader.num_channels)
        d_vis = torch.empty((nrow_total, nchan), dtype=torch.complex64, device=dev)
        d_effw = torch.empty((nrow_total, nchan), dtype=torch.float32, device=dev)
        ctx = Context.get(dev.index)
        if nrow_total and nchan:
            per_row = nchan * 4 * (8 + 4 + 1)
            rows = int(upload_rows or max(1, (256 << 20) // per_row))
            nblocks = min(nrow_total, max(1, -(-nrow_total // rows)))
            blocks = ms_reader.partition(nblocks, 1) if nblocks > 1 else [ms_reader]
            rows = max(int(b.num_data_rows) for b in blocks)
            copy_stream = torch.cuda.Stream(device=dev)
            stage = [dict(vis=torch.empty((rows, nchan, 4), dtype=torch.complex64, pin_memory=True),
                          flg=torch.empty((rows, nchan, 4), dtype=torch.uint8, pin_memory=True),
                          wgt=torch.empty((rows, nchan, 4), dtype=torch.float32, pin_memory=True),
                          d_vis=torch.empty((rows, nchan, 4), dtype=torch.complex64, device=dev),
                          d_flg=torch.empty((rows, nchan, 4), dtype=torch.uint8, device=dev),
                          d_wgt=torch.empty((rows, nchan, 4), dtype=torch.float32, device=dev),
                          sent=torch.cuda.Event(), used=torch.cuda.Event()) for _ in range(min(2, len(blocks)))]
            lib_stream = torch.cuda.ExternalStream(lib().cip_ctx_stream(ctx.handle), device=dev)
            row0 = 0
            for k, blk in enumerate(blocks):
                st = stage[k % len(stage)]
                n = int(blk.num_data_rows)
                vis = np.asarray(blk.visibilities())
                if vis.ndim != 3 or vis.shape[2] != 4:
                    raise ValueError("expected 4 polarisation products")
                st["used"].synchronize()  # the conversion that last read this pair of buffers has finished
                st["vis"][:n].numpy()[...] = vis
                st["flg"][:n].numpy()[...] = np.asarray(blk.flags()).view(np.uint8) if np.asarray(blk.flags()).dtype == np.bool_ \
                    else np.asarray(blk.flags(), dtype=np.uint8)
                st["wgt"][:n].numpy()[...] = blk.weights()
                with torch.cuda.stream(copy_stream):
                    st["d_vis"][:n].copy_(st["vis"][:n], non_blocking=True)
                    st["d_flg"][:n].copy_(st["flg"][:n], non_blocking=True)
                    st["d_wgt"][:n].copy_(st["wgt"][:n], non_blocking=True)
                    st["sent"].record(copy_stream)
                lib_stream.wait_event(st["sent"])
                with ctx.lock:  # asynchronous: the call returns once the kernel is enqueued behind the copy
                    check(lib().cip_stokes_i_async(ctx.handle, st["d_vis"].data_ptr(), st["d_flg"].data_ptr(),
                                                   st["d_wgt"].data_ptr(), n, nchan, d_vis[row0:row0 + n].data_ptr(),
                                                   d_effw[row0:row0 + n].data_ptr()))
                st["used"].record(lib_stream)
                row0 += n
            lib_stream.synchronize()
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


def invert_tiles(tiles, channel_frequencies, num_pixels: int, pixel_size_asec: float, *, epsilon: float = EPSILON,
                 accumulate_fp64: bool = False) -> NDArray:
    """Dirty image of a measurement set that was reordered into uvw tiles (``uvw_tiling.reorder_by_uvw_tile``):
    ``tiles`` are ``Tile`` objects or paths of tile chunk files; normalised by the number of samples (tiles carry
    Stokes-I visibilities without weights).  Same image arguments as ``invert_measurement_set``."""
    from pathlib import Path as _Path
    from .uvw_tiling import Tile
    from .wgridder import tiles2dirty
    loaded = [Tile.load_npz(t) if isinstance(t, (str, _Path)) else t for t in tiles]
    px = _pixel_size_lm(pixel_size_asec)
    image, info = tiles2dirty(loaded, channel_frequencies, num_pixels, num_pixels, px, px, epsilon=epsilon,
                              double_precision_accumulation=accumulate_fp64, return_info=True)
    total = np.float32(info["scan"]["sum_wgt"])
    return (1.0 / total) * image if total else image


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
        for rep in _grid_shards(ctx, [a for a, _sh, _w in args_list], plan, image, reports is not None):
            reports.append(rep)
    return image, plan, sumw


def _grid_shards(ctx: Context, args: list, plan: Plan, image, want_reports: bool = True) -> list[dict]:
    """Add the shards' uncorrected dirty images to ``image``.  Device-resident shards go through ONE call
    (cip_grid_accumulate_many: one plane stack, one set of FFTs for all of them); host shards one by one, each
    streamed in behind its own binning and gridding.  Stage reports are only requested when wanted: timing the
    stages of a streamed host shard makes the host wait for every chunk, i.e. gives up the overlap of upload and
    compute."""
    L = lib()
    reports = []
    if len(args) > 1 and all(a.memspace == 1 for a in args):
        arr = (_cabi.InvertArgs * len(args))(*args)
        rep = Report()
        with ctx.lock:
            check(L.cip_grid_accumulate_many(ctx.handle, arr, len(args), C.byref(plan), image.data_ptr(),
                                             C.byref(rep) if want_reports else None))
        return [rep.as_dict()] if want_reports else []
    for a in args:
        rep = Report()
        with ctx.lock:
            check(L.cip_grid_accumulate(ctx.handle, C.byref(a), C.byref(plan), image.data_ptr(),
                                        C.byref(rep) if want_reports else None))
        if want_reports:
            reports.append(rep.as_dict())
    return reports


def image_to_host(image) -> NDArray:
    """Device image -> numpy array through pinned memory (torch's caching pinned allocator keeps the block):
    a pageable read-back of a 16384^2 image runs at a fifth of the PCIe rate."""
    torch = _torch()
    if image.numel() * image.element_size() < (1 << 20):
        return image.cpu().numpy()
    host = torch.empty(image.shape, dtype=image.dtype, pin_memory=True)
    host.copy_(image, non_blocking=True)
    torch.cuda.current_stream(image.device).synchronize()
    return host.numpy()


# ---- the same, sharded by |w| (SURVEY.md 8e option 3; BASELINE.json configs[4]) --------------------------------

W_HIST_BINS = 4096


def w_slab_edges(counts, wmin: float, wmax: float, nslabs: int, *, plane_cost: float = 0.0, dw: float = 0.0,
                 bin_factor=None) -> list[float]:
    """Cut the |w| axis into ``nslabs`` slabs of (nearly) equal COST from a histogram of ``len(counts)`` equal bins
    over [wmin, wmax]: a bin costs its samples plus ``plane_cost`` sample-equivalents per w plane it spans (bin
    width / dw) -- with plane_cost = 0 the slabs hold equal numbers of samples, with a large one equal numbers of
    planes (what matters when the per-plane FFTs dominate).  Returns nslabs + 1 edges, the first 0 and the last
    +inf, so that every sample falls into exactly one slab [edges[s], edges[s + 1]).  ``bin_factor`` (one number per
    bin) scales the cost of the bins: what ``rebalanced_bin_factor`` has learnt from measured times."""
    cost = np.asarray(counts, dtype=np.float64).copy()
    if plane_cost > 0 and dw > 0 and wmax > wmin:
        cost += plane_cost * (wmax - wmin) / len(cost) / dw
    if bin_factor is not None:
        cost *= np.asarray(bin_factor, dtype=np.float64)
    cum = np.concatenate([[0.0], np.cumsum(cost)])
    total = cum[-1]
    edges = [0.0]
    for s in range(1, nslabs):
        # bin boundary at which the cumulative cost first reaches s / nslabs of the total
        b = int(np.searchsorted(cum, total * s / nslabs, side="left"))
        b = min(max(b, 1), len(cost))
        edges.append(max(wmin + (wmax - wmin) * b / len(cost), edges[-1]))
    edges.append(float("inf"))
    return edges


def rebalanced_bin_factor(bin_factor, edges, wmin: float, wmax: float, times, *, gain: float = 1.3,
                          done_within: float = 0.03):
    """One step of measured load balancing of the |w| slabs: ``times[s]`` is what slab s (cut at ``edges``) took on its
    rank; the cost of every histogram bin inside slab s is scaled by (times[s] / mean(times)) ** gain, so that the
    next cut gives slow slabs less and fast slabs more.  Every slab carries a fixed part (W - 1 extra planes), so its
    time answers less than proportionally to its content; the gain above 1 makes up for that without overshooting
    while that fixed part is at least a quarter of a slab's time.  Returns the new factors (the old ones when the
    slowest slab is within ``done_within`` of the mean).  Pure function of its arguments: every rank computes the
    same result from the gathered times."""
    f = np.asarray(bin_factor, dtype=np.float64).copy()
    t = np.asarray(times, dtype=np.float64)
    if len(t) < 2 or not np.all(t > 0) or t.max() <= (1.0 + done_within) * t.mean():
        return f
    nb = len(f)
    centres = wmin + (wmax - wmin) * (np.arange(nb) + 0.5) / nb
    slab = np.clip(np.searchsorted(np.asarray(edges[1:-1], dtype=np.float64), centres, side="right"), 0, len(t) - 1)
    f *= (t / t.mean())[slab] ** gain
    return f / f.mean()


# problem signature -> bin factors learnt over the calls so far (an imaging run inverts the same uvw many times)
_SLAB_BALANCE: dict = {}


# sample-equivalents of one w plane per grid cell (FFT + w-screen ~ 12 ps per cell, binning + gridding ~ 65 ps per
# sample on a B200: profiles/r2*_bench.json), for the cost model of the slabs
PLANE_COST_PER_CELL = 0.2


def _shard_args(sh: StokesIGridderInput, num_pixels, pixel_size_lm, epsilon, accumulate_fp64, w_slab=None):
    effw = sh.effective_weights()
    a = make_args(sh.uvw, sh.channel_frequencies, sh.visibilities, effw, None, num_pixels, num_pixels, pixel_size_lm,
                  pixel_size_lm, epsilon, True, accumulate_fp64, 0 if isinstance(sh.visibilities, np.ndarray) else 1,
                  w_slab=w_slab)
    return a, effw


def _exchange_rows_by_slab(shards, edges, *, group, device):
    """All-to-all of visibility rows over the process group: rank s receives, from every rank, the rows of every
    shard whose |w| range (|w_metres| x the shard's lowest .. highest frequency / c) touches slab s.  A row that
    straddles a slab edge goes to both neighbours; the exact sample-level split happens in the library (w_lo, w_hi).
    Returns the received shards (one per sending rank and shard of that rank)."""
    torch = _torch()
    dist = torch.distributed
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    c = 299792458.0
    meta = [None] * world  # every rank's list of channel-frequency arrays
    dist.all_gather_object(meta, [np.asarray(sh.channel_frequencies.cpu() if hasattr(sh.channel_frequencies, "cpu")
                                             else sh.channel_frequencies, dtype=np.float64) for sh in shards], group=group)
    received = []
    for k in range(max(len(m) for m in meta)):
        sh = shards[k] if k < len(shards) else None
        send_rows = []
        if sh is not None:
            aw = sh.uvw[:, 2].abs()
            f = meta[rank][k]
            lo, hi = aw * (float(f.min()) / c), aw * (float(f.max()) / c)
            for s in range(world):
                send_rows.append(torch.nonzero((hi >= edges[s]) & (lo < edges[s + 1])).reshape(-1))
        else:
            send_rows = [torch.zeros(0, dtype=torch.int64, device=device) for _ in range(world)]
        nsend = torch.tensor([len(r) for r in send_rows], dtype=torch.int64, device=device)
        nrecv = torch.empty_like(nsend)
        dist.all_to_all_single(nrecv, nsend, group=group)
        nsend_l, nrecv_l = nsend.tolist(), nrecv.tolist()
        order = torch.cat(send_rows) if sh is not None else torch.zeros(0, dtype=torch.int64, device=device)
        nchan_of = [len(meta[r][k]) if k < len(meta[r]) else 0 for r in range(world)]

        def swap(rows_tensor, width_of, dtype):
            """rows_tensor: (sum nsend, width(rank)) of dtype; receives (sum over r of nrecv[r] * width_of[r],)."""
            send = rows_tensor.reshape(-1).contiguous()
            w_mine = width_of[rank]
            out = torch.empty(sum(n * width_of[r] for r, n in enumerate(nrecv_l)), dtype=dtype, device=device)
            dist.all_to_all_single(out, send, [n * width_of[r] for r, n in enumerate(nrecv_l)],
                                   [n * w_mine for n in nsend_l], group=group)
            return out

        if sh is not None:
            uvw_s = sh.uvw.index_select(0, order)
            vis_s = torch.view_as_real(sh.visibilities.index_select(0, order))
            wgt_s = sh.effective_weights().index_select(0, order)
        else:
            uvw_s = torch.zeros((0, 3), dtype=torch.float64, device=device)
            vis_s = torch.zeros((0, 0, 2), dtype=torch.float32, device=device)
            wgt_s = torch.zeros((0, 0), dtype=torch.float32, device=device)
        uvw_r = swap(uvw_s, [3] * world, torch.float64)
        vis_r = swap(vis_s, [2 * n for n in nchan_of], torch.float32)
        wgt_r = swap(wgt_s, nchan_of, torch.float32)
        o3 = ov = ow = 0
        for r in range(world):
            n, nch = nrecv_l[r], nchan_of[r]
            if n and nch:
                freq = torch.from_numpy(meta[r][k]).to(device)
                received.append(StokesIGridderInput(
                    freq, None, uvw_r[o3:o3 + 3 * n].reshape(n, 3),
                    torch.view_as_complex(vis_r[ov:ov + 2 * n * nch].reshape(n, nch, 2)),
                    wgt_r[ow:ow + n * nch].reshape(n, nch)))
            o3 += 3 * n
            ov += 2 * n * nch
            ow += n * nch
    return received


def accumulate_shards_by_w(shards: Sequence[StokesIGridderInput], num_pixels: int, pixel_size_lm: float, *,
                           epsilon: float = EPSILON, accumulate_fp64: bool = False, group=None,
                           device: Optional[int] = None, nslabs: Optional[int] = None, reports: Optional[list] = None):
    """Grid this rank's share of a |w|-sharded invert into one uncorrected partial image.

    The ranks agree on the global |w| range and on (sigma, W, dw, nshift) as in ``accumulate_shards``, sum their |w|
    histograms and cut the axis into one slab of equal population per rank; the visibility rows are exchanged so that
    rank s holds what touches slab s (NCCL all-to-all -- the one exchange step of this sharding), and every rank
    grids its slab with a plan of ITS w range: about nplanes / world + W planes and FFTs instead of nplanes.  A single
    process (no group) walks ``nslabs`` slabs one after the other, which bounds the resident plane stack the same way.
    Returns (image CUDA tensor, Plan of the last slab gridded here, global sum of weights); the partial images of
    the ranks add up as in ``accumulate_shards`` and take the same ``reduce_and_correct``."""
    torch = _torch()
    dev = _device(device)
    ctx = Context.get(dev.index)
    L = lib()
    world = torch.distributed.get_world_size(group) if _dist_ready() else 1
    rank = torch.distributed.get_rank(group) if _dist_ready() else 0
    nslabs = world if world > 1 else max(int(nslabs or 1), 1)
    to_dev = lambda a, dt: a if not isinstance(a, np.ndarray) else torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    shards = [StokesIGridderInput(to_dev(sh.channel_frequencies, np.float64), None, to_dev(sh.uvw, np.float64),
                                  to_dev(sh.visibilities, np.complex64), to_dev(sh.effective_weights(), np.float32))
              for sh in shards]
    torch.cuda.current_stream(dev).synchronize()

    def scan_all(items, w_slab=None):
        wmin, wmax, nactive, sumw, keep = np.inf, 0.0, 0, 0.0, []
        for sh in items:
            a, effw = _shard_args(sh, num_pixels, pixel_size_lm, epsilon, accumulate_fp64, w_slab)
            scan = ScanResult()
            with ctx.lock:
                check(L.cip_scan(ctx.handle, C.byref(a), C.byref(scan)))
            keep.append((a, sh, effw))
            if scan.nactive:
                wmin, wmax = min(wmin, scan.wmin), max(wmax, scan.wmax)
            nactive += scan.nactive
            sumw += scan.sum_wgt
        return wmin, wmax, nactive, sumw, keep

    import time
    phases, t_last = {}, time.perf_counter()

    def mark(name):  # host clock between points where the device is idle anyway
        nonlocal t_last
        torch.cuda.current_stream(dev).synchronize()
        now = time.perf_counter()
        phases[name] = phases.get(name, 0.0) + 1e3 * (now - t_last)
        t_last = now

    wmin, wmax, nactive, sumw, keep = scan_all(shards)
    wmin, wmax, nactive, sumw = agree_on_shard_statistics(wmin, wmax, nactive, sumw, group=group, device=dev)
    mark("ms_scan")
    image = torch.zeros((num_pixels, num_pixels), dtype=torch.float64 if accumulate_fp64 else torch.float32, device=dev)
    if nactive == 0:
        plan = _cabi.plan_make(num_pixels, num_pixels, pixel_size_lm, pixel_size_lm, epsilon, 0.0, 0.0, 0,
                               do_wstacking=True, fp64=accumulate_fp64)
        return image, plan, sumw
    common = _cabi.plan_make(num_pixels, num_pixels, pixel_size_lm, pixel_size_lm, epsilon, wmin, wmax, nactive,
                             do_wstacking=True, fp64=accumulate_fp64, sigma=float(os.environ.get("CIP_SIGMA", 0) or 0),
                             support=int(os.environ.get("CIP_SUPPORT", 0) or 0))
    counts = np.zeros(W_HIST_BINS, dtype=np.int64)
    for a, _sh, _w in keep:
        with ctx.lock:
            check(L.cip_w_histogram(ctx.handle, C.byref(a), wmin, wmax, W_HIST_BINS, counts.ctypes.data))
    if world > 1:
        t = torch.from_numpy(counts).to(dev)
        torch.distributed.all_reduce(t, group=group)
        counts = t.cpu().numpy()
    # the slabs are cut by modelled cost, corrected by what the ranks measured on earlier calls of the same problem
    balance_key = (world, num_pixels, float(pixel_size_lm), float(epsilon), bool(accumulate_fp64), int(nactive),
                   float(wmin), float(wmax))
    bin_factor = _SLAB_BALANCE.get(balance_key) if world > 1 and not os.environ.get("CIP_NO_SLAB_REBALANCE") else None
    edges = w_slab_edges(counts, wmin, wmax, nslabs, plane_cost=PLANE_COST_PER_CELL * common.nu * common.nv, dw=common.dw,
                         bin_factor=bin_factor)
    mark("ms_histogram")
    mine = shards
    if world > 1:
        mine = _exchange_rows_by_slab(shards, edges, group=group, device=dev)
        mark("ms_exchange")
    plan = common
    t_own = time.perf_counter()  # the device is idle here (mark synchronises)
    for s in ([rank] if world > 1 else range(nslabs)):
        slab = (edges[s], min(edges[s + 1], 1e300))
        smin, smax, sact, _ssum, skeep = scan_all(mine, slab)
        mark("ms_slab_scan")
        if sact == 0:
            continue
        plan = _cabi.plan_make(num_pixels, num_pixels, pixel_size_lm, pixel_size_lm, epsilon, smin, smax, sact,
                               do_wstacking=True, fp64=accumulate_fp64, sigma=common.sigma, support=common.support)
        assert (plan.nu, plan.nv, plan.dw, plan.nshift) == (common.nu, common.nv, common.dw, common.nshift)
        for rep in _grid_shards(ctx, [a for a, _sh, _w in skeep], plan, image, reports is not None):
            reports.append(dict(rep, nplanes=plan.nplanes, slab=s))
        mark("ms_grid_call")
    if world > 1 and not os.environ.get("CIP_NO_SLAB_REBALANCE"):
        # every rank's time for its slab -> the bin factors of the next call (one small all-gather)
        mine = torch.tensor([time.perf_counter() - t_own], dtype=torch.float64, device=dev)
        every = torch.empty(world, dtype=torch.float64, device=dev)
        torch.distributed.all_gather_into_tensor(every, mine, group=group)
        if len(_SLAB_BALANCE) >= 64 and balance_key not in _SLAB_BALANCE:
            _SLAB_BALANCE.clear()  # a long-running process that meets many different problems: start over
        _SLAB_BALANCE[balance_key] = rebalanced_bin_factor(np.ones(W_HIST_BINS) if bin_factor is None else bin_factor,
                                                           edges, wmin, wmax, every.cpu().numpy())
        mark("ms_balance")
    if reports:
        reports[-1]["phases_ms"] = phases
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
    return image_to_host(image)
