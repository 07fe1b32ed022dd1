"""Drop-in for ``ska_sdp_cip.uvw_tiling.tiling_plan`` (reference:
src/ska_sdp_cip/uvw_tiling/tiling_plan.py): same names, arguments and result types; the binning
itself runs on the GPU through ``cip_tile_bin`` (include/cip.h) and is bit-exact with the
reference's tile indices and orderings.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import NamedTuple, Optional

import numpy as np
from numpy.typing import NDArray

from .. import _cabi
from .._cabi import Context, check, lib

TileCoords = tuple[int, int, int]
"""Tile index of the form (iu, iv, iw)"""


class RowSliceId(NamedTuple):
    """A slice of one visibility row along the frequency axis (tiling_plan.py:16-23)."""

    irow: int
    chan_start: int
    chan_stop: int


TileMapping = dict[TileCoords, list[RowSliceId]]


class DeviceTileMapping:
    """The binning result kept on the GPU (handle of ``cip_tile_result``) plus lazily fetched
    host arrays.  ``to_mapping()`` materialises the reference's dict."""

    def __init__(self, ctx: Context, handle, n_tiles: int, n_slices: int, nrow: int, nchan: int):
        self._ctx, self._h = ctx, handle
        self.n_tiles, self.n_slices, self.nrow, self.nchan = n_tiles, n_slices, nrow, nchan
        self._arrays = None

    def arrays(self):
        """(tile_coords (nt,3) i64, tile_offsets (nt+1,) i64, irow i64, chan_start i32, chan_stop i32)."""
        if self._arrays is None:
            coords = np.empty((self.n_tiles, 3), dtype=np.int64)
            offsets = np.empty(self.n_tiles + 1, dtype=np.int64)
            irow = np.empty(self.n_slices, dtype=np.int64)
            c0 = np.empty(self.n_slices, dtype=np.int32)
            c1 = np.empty(self.n_slices, dtype=np.int32)
            with self._ctx.lock:
                check(lib().cip_tile_fetch(self._ctx.handle, self._h, coords.ctypes.data, offsets.ctypes.data,
                                           irow.ctypes.data, c0.ctypes.data, c1.ctypes.data))
            self._arrays = (coords, offsets, irow, c0, c1)
        return self._arrays

    def to_mapping(self) -> TileMapping:
        coords, offsets, irow, c0, c1 = self.arrays()
        irow_l, c0_l, c1_l = irow.tolist(), c0.tolist(), c1.tolist()  # python ints, as the reference
        off = offsets.tolist()
        out: TileMapping = {}
        for t, key in enumerate(map(tuple, coords.tolist())):
            a, b = off[t], off[t + 1]
            out[key] = list(map(RowSliceId, irow_l[a:b], c0_l[a:b], c1_l[a:b]))
        return out

    def gather(self, vis: NDArray, uvw: NDArray):
        """All tiles' data at once, tile-contiguous (the scatter half of the counting sort,
        tile.py:83-115): (visibilities (nvis,) c64, vis_offsets (n_slices+1,) i64, uvw (n_slices,3) f64)."""
        vis = np.ascontiguousarray(vis, dtype=np.complex64)
        uvw = np.ascontiguousarray(uvw, dtype=np.float64)
        if vis.shape != (self.nrow, self.nchan):
            raise ValueError(f"vis must have shape {(self.nrow, self.nchan)}")
        vout = np.empty(self.nrow * self.nchan, dtype=np.complex64)
        voff = np.empty(self.n_slices + 1, dtype=np.int64)
        uout = np.empty((self.n_slices, 3), dtype=np.float64)
        with self._ctx.lock:
            check(lib().cip_tile_gather(self._ctx.handle, self._h, vis.ctypes.data, uvw.ctypes.data, self.nrow,
                                        self.nchan, 0, vout.ctypes.data, voff.ctypes.data, uout.ctypes.data))
        return vout, voff, uout

    def close(self):
        if self._h:
            lib().cip_tile_free(self._ctx.handle, self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


def device_tile_mapping(uvw: NDArray, tile_size: tuple[float, float, float], channel_freqs: NDArray, *,
                        row_offset: int = 0, device: int = -1) -> DeviceTileMapping:
    """Run the GPU binning and keep the result on the device."""
    uvw = np.ascontiguousarray(uvw, dtype=np.float64)
    if uvw.ndim != 2 or uvw.shape[1] != 3:
        if uvw.size == 0:
            uvw = uvw.reshape(0, 3)
        else:
            raise ValueError(f"uvw must have shape (nrow, 3), got {uvw.shape}")
    freqs = np.ascontiguousarray(channel_freqs, dtype=np.float64).reshape(-1)
    if len(uvw) and len(freqs) == 0:
        raise IndexError("index 0 is out of bounds for axis 0 with size 0")  # what the reference raises
    ts = (C.c_double * 3)(*[float(x) for x in tile_size])
    ctx = Context.get(device)
    h, nt, ns = C.c_void_p(), C.c_int64(), C.c_int64()
    with ctx.lock:
        check(lib().cip_tile_bin(ctx.handle, uvw.ctypes.data, len(uvw), freqs.ctypes.data, len(freqs), C.byref(ts),
                                 int(row_offset), 0, C.byref(h), C.byref(nt), C.byref(ns)))
    return DeviceTileMapping(ctx, h, nt.value, ns.value, len(uvw), len(freqs))


def create_uvw_tile_mapping_sequential(uvw: NDArray, tile_size: tuple[float, float, float],
                                       channel_freqs: NDArray, *, row_offset: int = 0) -> TileMapping:
    """Bin the UVW coordinates of visibilities by UVW tile (tiling_plan.py:29-61)."""
    dm = device_tile_mapping(uvw, tile_size, channel_freqs, row_offset=row_offset)
    try:
        return dm.to_mapping()
    finally:
        dm.close()


def create_uvw_tile_mapping(uvw: NDArray, tile_size: tuple[float, float, float], channel_freqs: NDArray, *,
                            processes: Optional[int] = os.cpu_count()) -> TileMapping:
    """Bin the UVW coordinates of visibilities by UVW tile (tiling_plan.py:84-134).

    Keys are integer 3-tuples in first-appearance order, values lists of ``RowSliceId`` in ascending
    (irow, chan_start) order.  ``processes`` is accepted for compatibility and ignored: the row-block
    fan-out and merge of the reference (tiling_plan.py:114-147) is one GPU pass here and yields the
    same dict.  ``channel_freqs`` must be monotonic, as in the reference.
    """
    del processes
    return create_uvw_tile_mapping_sequential(uvw, tile_size, channel_freqs)


def merge_tile_mappings(tile_mappings: list[TileMapping]) -> TileMapping:
    """Merge tile mappings into one, keeping first-appearance key order (tiling_plan.py:137-147)."""
    merged: TileMapping = {}
    for mapping in tile_mappings:
        for coords, slices in mapping.items():
            merged.setdefault(coords, []).extend(slices)
    return merged
