"""Drop-in for ``ska_sdp_cip.uvw_tiling.reorder`` (reference: src/ska_sdp_cip/uvw_tiling/reorder.py:19-205):
convert to Stokes I and reorder the visibilities into UVW tile chunk files, in two passes --

1. per time interval: bin the interval by tile (``cip_tile_bin`` on the GPU, bit-exact with the reference's
   ``create_uvw_tile_mapping``), gather every tile's visibilities with one device gather
   (``cip_tile_gather``, replacing the per-slice Python loop of ``reorder_time_interval``, reorder.py:129-155)
   and write one npz per (tile, interval);
2. per tile: rechunk its interval files into files of at most ``max_vis_per_chunk`` visibilities.

File names and formats are the reference's (reorder.py:168-171, 186-192).  With a dask ``client`` the two
passes fan out as tasks exactly like the reference (each worker needs a GPU for pass 1); with ``client=None``
they run in this process.
"""
from __future__ import annotations

import itertools
from pathlib import Path
from typing import Optional

import numpy as np
from numpy.typing import NDArray

from .tile import Tile, rechunk_tiles_on_disk, tiles_from_device_mapping
from .tiling_plan import TileCoords, TileMapping, create_uvw_tile_mapping, device_tile_mapping


def _tile_basename(tile_coords: TileCoords) -> str:
    u, v, w = tile_coords
    return f"tile_iu{u:+03d}_iv{v:+03d}_iw{w:+03d}"


def _tile_filename(tile_coords: TileCoords, interval_index: int) -> str:
    return f"{_tile_basename(tile_coords)}_interval{interval_index:02d}.npz"


def stokes_i_visibilities(vis4: NDArray) -> NDArray:
    """``0.5 * (v0 + v3)`` in the reference's operation order and dtype (reorder.py:145)."""
    return 0.5 * (vis4[..., 0] + vis4[..., 3])


def create_time_interval_tile_mapping(ms_reader, tile_size: TileCoords, channel_freqs: NDArray) -> TileMapping:
    """The UVW tile mapping of one time interval (reorder.py:114-126)."""
    return create_uvw_tile_mapping(ms_reader.uvw(), tile_size, channel_freqs)


def reorder_time_interval(ms_reader, tile_mapping: Optional[TileMapping], outdir: Path, *, interval_index: int,
                          tile_size: Optional[TileCoords] = None,
                          channel_freqs: Optional[NDArray] = None) -> list[TileCoords]:
    """Write one npz per tile present in the interval and return their coordinates (reorder.py:129-155).

    With a pre-computed ``tile_mapping`` (the reference's calling convention) the tiles are cut on the host;
    with ``tile_mapping=None`` and ``tile_size`` / ``channel_freqs`` given, binning and gather both run on the
    GPU and the dict of Python objects is never built.
    """
    uvw = ms_reader.uvw()
    stokes_i = stokes_i_visibilities(ms_reader.visibilities())
    if tile_mapping is None:
        dm = device_tile_mapping(uvw, tile_size, channel_freqs)
        try:
            tiles = tiles_from_device_mapping(dm, stokes_i, uvw)
        finally:
            dm.close()
    else:
        tiles = [Tile._from_jagged_visibilities_slice(stokes_i, uvw, coords, row_slices)  # noqa: SLF001
                 for coords, row_slices in tile_mapping.items()]
    for tile in tiles:
        tile.save_npz(outdir / _tile_filename(tile.coords, interval_index))
    return [tile.coords for tile in tiles]


def rechunk_tile_chunk_group(tile_coords: TileCoords, outdir: Path, *, max_vis_per_chunk: int = 5_000_000
                             ) -> list[Path]:
    """Rechunk the interval files of one tile inside ``outdir`` and delete them (reorder.py:158-183)."""
    base = _tile_basename(tile_coords)
    inputs = sorted(outdir.glob(f"{base}_interval*.npz"))
    outputs = rechunk_tiles_on_disk(inputs, outdir, base, max_vis_per_chunk=max_vis_per_chunk)
    for path in inputs:
        path.unlink()
    return outputs


def reorder_by_uvw_tile(ms_reader, tile_size: TileCoords, outdir: Path, client=None, *,
                        num_time_intervals: Optional[int] = None, max_vis_per_chunk: int = 5_000_000
                        ) -> list[Path]:
    """Reorder the measurement set into UVW tile chunk files under ``outdir``; returns their paths
    (reorder.py:19-111; same arguments, ``client`` may additionally be None for in-process execution)."""
    if num_time_intervals is None:
        nworkers = len(client.scheduler_info()["workers"]) if client is not None else 1
        num_time_intervals = max(2 * nworkers, 2)
    outdir = Path(outdir).resolve()
    channel_freqs = np.asarray(ms_reader.channel_frequencies())
    intervals = ms_reader.partition(num_time_intervals, 1)

    if client is None:
        present: set[TileCoords] = set()
        for index, reader in enumerate(intervals):
            present.update(reorder_time_interval(reader, None, outdir, interval_index=index, tile_size=tile_size,
                                                 channel_freqs=channel_freqs))
        return list(itertools.chain.from_iterable(
            rechunk_tile_chunk_group(coords, outdir, max_vis_per_chunk=max_vis_per_chunk)
            for coords in sorted(present)))

    coords_futures = []
    for index, reader in enumerate(intervals):
        mapping = client.submit(create_time_interval_tile_mapping, reader, tile_size, channel_freqs,
                                resources={"processing_slots": 1})
        coords_futures.append(client.submit(reorder_time_interval, reader, mapping, outdir, interval_index=index))
    present = set()
    for future in coords_futures:
        present.update(future.result())
    rechunk_futures = [client.submit(rechunk_tile_chunk_group, coords, outdir, max_vis_per_chunk=max_vis_per_chunk)
                       for coords in present]
    return list(itertools.chain.from_iterable(f.result() for f in rechunk_futures))
