"""Drop-in for ``ska_sdp_cip.uvw_tiling.tile`` (reference: src/ska_sdp_cip/uvw_tiling/tile.py): the ``Tile``
container (:14-124: same fields, npz format, constructor helpers), ``concatenate_tiles`` (:127-152),
``split_tile`` (:155-213) and ``rechunk_tiles_on_disk`` (:216-265).  ``tiles_from_device_mapping`` builds all
tiles of a block from one GPU gather (cip_tile_gather) instead of a Python loop per slice."""
from __future__ import annotations

import os
from dataclasses import dataclass
from pathlib import Path
from typing import Iterable, Sequence, Union

import numpy as np
from numpy.typing import NDArray

from .tiling_plan import DeviceTileMapping, RowSliceId, TileCoords


@dataclass(repr=False)
class Tile:
    """Visibility data and metadata of one UVW tile."""

    coords: TileCoords
    uvw: NDArray
    visibilities: NDArray
    channel_start_indices: NDArray
    channel_stop_indices: NDArray

    @property
    def num_rows(self) -> int:
        return len(self.uvw)

    @property
    def num_visibilities(self) -> int:
        return len(self.visibilities)

    def save_npz(self, path: Union[str, os.PathLike]) -> None:
        """Same keys and dtypes as the reference (tile.py:40-51)."""
        np.savez(path, coords=np.asarray(self.coords).astype(int), uvw=self.uvw, visibilities=self.visibilities,
                 channel_start_indices=self.channel_start_indices, channel_stop_indices=self.channel_stop_indices)

    @classmethod
    def load_npz(cls, path: Union[str, os.PathLike]) -> "Tile":
        with np.load(path) as npz:
            return cls(tuple(map(int, npz["coords"])), npz["uvw"], npz["visibilities"],
                       npz["channel_start_indices"], npz["channel_stop_indices"])

    @classmethod
    def _zeros(cls, coords: TileCoords, num_row_slices: int, num_vis: int) -> "Tile":
        return cls(coords, np.zeros((num_row_slices, 3), dtype=float), np.zeros(num_vis, dtype=np.complex64),
                   np.zeros(num_row_slices, dtype=int), np.zeros(num_row_slices, dtype=int))

    @classmethod
    def _from_jagged_visibilities_slice(cls, vis: NDArray, uvw: NDArray, coords: TileCoords,
                                        row_slices: list[RowSliceId]) -> "Tile":
        """Extract the given row slices from a (row, freq) block (tile.py:83-115).  Host-side
        convenience for a single tile; whole blocks go through ``tiles_from_device_mapping``."""
        irow = np.fromiter((r.irow for r in row_slices), dtype=np.int64, count=len(row_slices))
        c0 = np.fromiter((r.chan_start for r in row_slices), dtype=np.int64, count=len(row_slices))
        c1 = np.fromiter((r.chan_stop for r in row_slices), dtype=np.int64, count=len(row_slices))
        parts = [vis[r, a:b] for r, a, b in zip(irow.tolist(), c0.tolist(), c1.tolist())]
        flat = np.concatenate(parts).astype(np.complex64) if parts else np.zeros(0, dtype=np.complex64)
        return cls(coords, np.asarray(uvw, dtype=float)[irow].reshape(-1, 3), flat, c0, c1)

    def __str__(self) -> str:
        return f"Tile(coords={self.coords}, nrows={self.num_rows}, nvis={self.num_visibilities})"

    __repr__ = __str__


def tiles_from_device_mapping(dm: DeviceTileMapping, vis: NDArray, uvw: NDArray) -> list[Tile]:
    """All tiles of a binned block, in the mapping's (first-appearance) order, from one GPU gather."""
    coords, offsets, _irow, c0, c1 = dm.arrays()
    flat, voff, suvw = dm.gather(vis, uvw)
    tiles = []
    for t in range(dm.n_tiles):
        a, b = int(offsets[t]), int(offsets[t + 1])
        tiles.append(Tile(tuple(int(x) for x in coords[t]), suvw[a:b].copy(), flat[voff[a]:voff[b]].copy(),
                          c0[a:b].astype(int), c1[a:b].astype(int)))
    return tiles


_ARRAY_FIELDS = ("uvw", "visibilities", "channel_start_indices", "channel_stop_indices")


def concatenate_tiles(tiles: Sequence[Tile]) -> Tile:
    """One tile holding the rows of all ``tiles`` in order (tile.py:127-152); they must share coordinates."""
    if not tiles:
        raise ValueError("Cannot concatenate empty sequence of tiles")
    coords = tiles[0].coords
    if any(t.coords != coords for t in tiles):
        raise ValueError("Cannot merge tiles with different coordinates")
    return Tile(coords, *(np.concatenate([getattr(t, name) for t in tiles]) for name in _ARRAY_FIELDS))


def split_tile(tile: Tile, max_vis_per_chunk: int) -> list[Tile]:
    """Cut a tile into consecutive chunks of at most ``max_vis_per_chunk`` visibilities (tile.py:155-213).
    Row slices are never split, so a single slice longer than the limit becomes a chunk of its own; the
    chunks are views of the tile's arrays.  Greedy like the reference: a chunk is closed when the next row
    slice would take it over the limit -- found by bisection on the cumulative sizes instead of a row loop."""
    sizes = (np.asarray(tile.channel_stop_indices, dtype=np.int64)
             - np.asarray(tile.channel_start_indices, dtype=np.int64))
    ends = np.cumsum(sizes)  # ends[i] = visibilities in rows 0..i
    nrows = len(sizes)
    chunks: list[Tile] = []
    row = 0
    while row < nrows:
        done = int(ends[row - 1]) if row else 0
        # last row whose cumulative count still fits; at least one row per chunk
        stop = max(int(np.searchsorted(ends, done + max_vis_per_chunk, side="right")), row + 1)
        vis_stop = int(ends[stop - 1])
        chunks.append(Tile(tile.coords, tile.uvw[row:stop], tile.visibilities[done:vis_stop],
                           tile.channel_start_indices[row:stop], tile.channel_stop_indices[row:stop]))
        row = stop
    return chunks


def rechunk_tiles_on_disk(tile_paths: Iterable[Path], outdir: Path, basename: str, *,
                          max_vis_per_chunk: int = 5_000_000) -> list[Path]:
    """Rewrite the tile chunk files ``tile_paths`` (same tile coordinates) as files named
    ``{basename}_chunkNNN.npz`` that are as full as possible without exceeding ``max_vis_per_chunk``
    visibilities; returns the written paths in order (tile.py:216-265)."""
    written: list[Path] = []

    def write(tile: Tile) -> None:
        path = outdir / f"{basename}_chunk{len(written):03d}.npz"
        tile.save_npz(path)
        written.append(path)

    pending: list[Tile] = []   # read but not yet written: together below the limit, or a single remainder
    for path in tile_paths:
        pending.append(Tile.load_npz(path))
        if sum(t.num_visibilities for t in pending) > max_vis_per_chunk:
            *full, rest = split_tile(concatenate_tiles(pending) if len(pending) > 1 else pending[0],
                                     max_vis_per_chunk)
            for chunk in full:
                write(chunk)
            pending = [rest]
    if pending:
        write(concatenate_tiles(pending) if len(pending) > 1 else pending[0])
    return written
