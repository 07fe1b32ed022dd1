"""Drop-in for the ``Tile`` container of the reference (src/ska_sdp_cip/uvw_tiling/tile.py:14-124):
same fields, npz format and constructor helpers.  ``tiles_from_device_mapping`` builds all tiles
of a block from one GPU gather (cip_tile_gather) instead of a Python loop per slice."""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Union

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
