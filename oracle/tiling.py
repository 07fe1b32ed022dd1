"""ORACLE (test infrastructure): numpy restatement of the reference's uvw tile binning.

Follows src/ska_sdp_cip/uvw_tiling/tiling_plan.py:29-61 (index formula and per-row loop),
:150-181 (maximal constant runs along the channel axis; the reference finds them by recursive
bisection, which equals plain run-length encoding for monotonic frequencies), :137-147 (dict
order = first appearance, list order = ascending (irow, chan_start)) and
src/ska_sdp_cip/uvw_tiling/tile.py:83-115 (gather of row slices).
PINNED: checked bit for bit against fixtures produced by the reference's own code
(tools/make_golden.py -> tests/golden/tiling_*.npz).
"""
from __future__ import annotations

import numpy as np

SPEED_OF_LIGHT = 299792458.0


def tile_indices(uvw, tile_size, channel_freqs):
    """(nrow, nchan, 3) int64: floor((freq/c) * (uvw/tile) + 0.5), three separately rounded float64 ops."""
    wavelength_inv = np.asarray(channel_freqs, dtype=np.float64).reshape(1, -1, 1) / SPEED_OF_LIGHT
    scaled = np.asarray(uvw, dtype=np.float64) / np.asarray(tile_size, dtype=np.float64)
    return np.floor(wavelength_inv * scaled[:, None, :] + 0.5).astype(np.int64)


def tile_mapping_arrays(uvw, tile_size, channel_freqs, row_offset=0):
    """Flat form of the mapping: (tile_coords (nt,3), tile_offsets (nt+1,), irow, ch_start, ch_stop),
    tiles in first-appearance order, slices grouped by tile in ascending (irow, ch_start) order."""
    uvw = np.asarray(uvw, dtype=np.float64).reshape(-1, 3)
    nrow = len(uvw)
    nchan = len(channel_freqs)
    if nrow == 0:
        z = np.zeros(0, dtype=np.int64)
        return np.zeros((0, 3), dtype=np.int64), np.zeros(1, dtype=np.int64), z, z.astype(np.int32), z.astype(np.int32)
    if nchan == 0:
        raise IndexError("channel_freqs is empty")
    idx = tile_indices(uvw, tile_size, channel_freqs)
    start = np.ones((nrow, nchan), dtype=bool)
    start[:, 1:] = np.any(idx[:, 1:] != idx[:, :-1], axis=2)
    rows, ch0 = np.nonzero(start)                      # (row, channel) scan order
    coords = idx[rows, ch0]
    nxt = np.empty_like(ch0)
    nxt[:-1] = np.where(rows[1:] == rows[:-1], ch0[1:], nchan)
    nxt[-1] = nchan
    uniq, first, inv = np.unique(coords, axis=0, return_index=True, return_inverse=True)
    inv = inv.reshape(-1)
    order = np.argsort(first, kind="stable")           # tiles by first appearance
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    tile_rank = rank[inv]
    perm = np.argsort(tile_rank, kind="stable")        # slices grouped by tile, scan order kept
    counts = np.bincount(tile_rank, minlength=len(uniq))
    offsets = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    return (uniq[order].astype(np.int64), offsets, (rows[perm] + row_offset).astype(np.int64),
            ch0[perm].astype(np.int32), nxt[perm].astype(np.int32))


def tile_mapping(uvw, tile_size, channel_freqs, row_offset=0):
    """dict[(iu,iv,iw)] -> list[(irow, chan_start, chan_stop)] with python ints, reference ordering."""
    coords, offsets, irow, c0, c1 = tile_mapping_arrays(uvw, tile_size, channel_freqs, row_offset)
    out = {}
    for t in range(len(coords)):
        key = tuple(int(x) for x in coords[t])
        s = slice(offsets[t], offsets[t + 1])
        out[key] = list(zip(irow[s].tolist(), c0[s].tolist(), c1[s].tolist()))
    return out


def gather_tile(vis, uvw, row_slices):
    """tile.py:83-115: (uvw (ns,3) f64, visibilities (nv,) c64, start (ns,) i64, stop (ns,) i64)."""
    ns = len(row_slices)
    nvis = sum(b - a for _, a, b in row_slices)
    t_uvw = np.zeros((ns, 3), dtype=float)
    t_vis = np.zeros(nvis, dtype=np.complex64)
    t_a = np.zeros(ns, dtype=int)
    t_b = np.zeros(ns, dtype=int)
    k = 0
    for i, (irow, a, b) in enumerate(row_slices):
        t_vis[k:k + b - a] = vis[irow, a:b]
        t_uvw[i] = uvw[irow]
        t_a[i], t_b[i] = a, b
        k += b - a
    return t_uvw, t_vis, t_a, t_b
