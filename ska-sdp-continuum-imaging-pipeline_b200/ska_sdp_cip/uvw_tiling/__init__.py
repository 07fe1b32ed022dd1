"""UVW tiling partition API (reference surface: src/ska_sdp_cip/uvw_tiling/__init__.py:1-17).
``reorder_by_uvw_tile`` (dask/disk orchestration) is out of the hot-path scope."""
from .tile import Tile, tiles_from_device_mapping
from .tiling_plan import (
    DeviceTileMapping,
    RowSliceId,
    TileCoords,
    TileMapping,
    create_uvw_tile_mapping,
    create_uvw_tile_mapping_sequential,
    device_tile_mapping,
    merge_tile_mappings,
)

__all__ = [
    "create_uvw_tile_mapping",
    "RowSliceId",
    "TileCoords",
    "TileMapping",
    "Tile",
]
