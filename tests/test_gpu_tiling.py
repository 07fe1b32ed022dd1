"""Bit-exact parity of the CUDA tile binning (cip_tile_bin / cip_tile_fetch / cip_tile_gather through
the C-ABI) with the reference's own uvw_tiling results (golden fixtures made by importing the
reference, tools/make_golden.py) and with the oracle on seeded inputs; at BASELINE.json configs[2]
scale through size-independent properties."""
import numpy as np
import pytest

from oracle import tiling

pytestmark = pytest.mark.gpu

CASES = ["meerkat_l", "small_tiles", "one_channel", "descending_freq", "half_integers"]
NAMES = ("coords", "offsets", "irow", "chan_start", "chan_stop")


@pytest.fixture(scope="module")
def T(built_lib):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from ska_sdp_cip import uvw_tiling
    return uvw_tiling


@pytest.mark.parametrize("case", CASES)
def test_tile_binning_matches_reference_golden(T, golden, case):
    g = golden(f"tiling_{case}")
    dm = T.device_tile_mapping(g["uvw"], tuple(g["tile_size"]), g["freq"], row_offset=int(g["row_offset"]))
    for name, got in zip(NAMES, dm.arrays()):
        assert got.dtype == g[name].dtype or np.can_cast(got.dtype, g[name].dtype), name
        assert np.array_equal(got, g[name]), name


@pytest.mark.parametrize("case", CASES)
def test_tile_gather_matches_reference_golden(T, golden, case):
    g = golden(f"tiling_{case}")
    off = int(g["row_offset"])
    uvw, freq = g["uvw"], g["freq"]
    rng = np.random.default_rng(int(g["vis_seed"]))
    vis = (rng.normal(size=(len(uvw), len(freq))) + 1j * rng.normal(size=(len(uvw), len(freq)))).astype(np.complex64)
    dm = T.device_tile_mapping(uvw, tuple(g["tile_size"]), freq, row_offset=off)
    tiles = T.tiles_from_device_mapping(dm, vis, uvw)
    assert len(tiles) == dm.n_tiles
    for t in range(int(g["n_gather_tiles"])):
        assert np.array_equal(tiles[t].uvw, g[f"tile{t}_uvw"])
        assert np.array_equal(tiles[t].visibilities, g[f"tile{t}_vis"])
        assert np.array_equal(tiles[t].channel_start_indices, g[f"tile{t}_start"])
        assert np.array_equal(tiles[t].channel_stop_indices, g[f"tile{t}_stop"])
    assert sum(t.num_visibilities for t in tiles) == vis.size


def test_mapping_has_reference_types_and_order(T, golden):
    g = golden("tiling_small_tiles")
    m = T.create_uvw_tile_mapping(g["uvw"], tuple(g["tile_size"]), g["freq"])
    want = tiling.tile_mapping(g["uvw"], tuple(g["tile_size"]), g["freq"])
    assert list(m) == list(want)                      # first-appearance key order
    for key, slices in m.items():
        assert all(type(k) is int for k in key)       # python ints (tiling_plan.py:163-165)
        assert all(isinstance(s, T.RowSliceId) and type(s.irow) is int for s in slices[:3])
        assert [tuple(s) for s in slices] == [tuple(s) for s in want[key]]


def test_every_sample_in_exactly_one_tile(T):
    """The reference's own test (tests/uvw_tiling/test_uvw_tiling_plan.py:7-32) on a MeerKAT-L-shaped layout."""
    rng = np.random.default_rng(0)
    nrow, nchan = 5000, 256
    uvw = rng.normal(size=(nrow, 3)) * np.array([4000, 4000, 300.0])
    uvw[::50] = 0  # autocorrelations
    freq = 856e6 + np.arange(nchan) * 214e6 / 256
    m = T.create_uvw_tile_mapping(uvw, (3000.0, 3000.0, 6000.0), freq)
    hit = np.zeros((nrow, nchan), dtype=int)
    for slices in m.values():
        for s in slices:
            hit[s.irow, s.chan_start:s.chan_stop] += 1
    assert (hit == 1).sum() == hit.size


@pytest.mark.parametrize("nrow,nchan,ts,seed", [(3000, 64, (300.0, 300.0, 600.0), 0),
                                                 (20000, 256, (50.0, 50.0, 100.0), 1),
                                                 (1, 1, (3000.0, 3000.0, 6000.0), 2),
                                                 (7, 1000, (10.0, 10.0, 10.0), 3),
                                                 (200000, 64, (3000.0, 3000.0, 6000.0), 4),
                                                 (5000, 33, (1e-2, 1e9, 1e9), 5)])
def test_random_layouts_match_oracle(T, nrow, nchan, ts, seed):
    rng = np.random.default_rng(seed)
    uvw = rng.normal(size=(nrow, 3)) * np.array([4000, 4000, 300.0])
    freq = 856e6 + np.arange(nchan) * 214e6 / 256
    dm = T.device_tile_mapping(uvw, ts, freq, row_offset=5)
    want = tiling.tile_mapping_arrays(uvw, ts, freq, row_offset=5)
    for name, got, exp in zip(NAMES, dm.arrays(), want):
        assert np.array_equal(got, exp), name


def test_half_integer_boundaries_round_like_numpy(T):
    """floor(x + 0.5) with x exactly on a boundary, negative zero, denormals and huge coordinates."""
    freq = np.array([299792458.0, 2 * 299792458.0])  # freq/c = 1, 2 exactly
    vals = np.array([0.5, -0.5, 1.5, -1.5, 2.5, 0.49999999999999994, -0.0, 5e-324, 2.0 ** 19 + 0.5, -2.0 ** 19 - 0.5,
                     1048575.5, 0.25, -0.25, 0.75, -0.75])
    uvw = np.stack([vals, vals[::-1], np.zeros_like(vals)], axis=1)
    dm = T.device_tile_mapping(uvw, (1.0, 1.0, 1.0), freq)
    want = tiling.tile_mapping_arrays(uvw, (1.0, 1.0, 1.0), freq)
    for name, got, exp in zip(NAMES, dm.arrays(), want):
        assert np.array_equal(got, exp), name


def test_tile_index_range_beyond_the_sort_key_is_refused(T):
    """The three tile indices of all slices are packed into one 64-bit sort key; a layout whose index
    extents need more bits is refused with ValueError (documented limit), never binned wrongly."""
    freq = np.array([299792458.0])
    uvw = np.array([[1e15, -1e15, 0.0], [-1e15, 1e15, 0.0]])
    with pytest.raises(ValueError, match="too large"):
        T.create_uvw_tile_mapping(uvw, (1.0, 1.0, 1.0), freq)


def test_empty_and_degenerate_inputs(T):
    freq = np.array([1e9, 1.1e9])
    m = T.create_uvw_tile_mapping(np.zeros((0, 3)), (1.0, 1.0, 1.0), freq)
    assert m == {}
    m = T.create_uvw_tile_mapping(np.zeros((3, 3)), (1.0, 1.0, 1.0), freq)
    assert m == {(0, 0, 0): [T.RowSliceId(0, 0, 2), T.RowSliceId(1, 0, 2), T.RowSliceId(2, 0, 2)]}
    with pytest.raises(IndexError):  # the reference indexes channel 0 of an empty frequency axis
        T.create_uvw_tile_mapping(np.zeros((3, 3)), (1.0, 1.0, 1.0), np.zeros(0))
    with pytest.raises(ValueError):
        T.create_uvw_tile_mapping(np.zeros((3, 2)), (1.0, 1.0, 1.0), freq)
    with pytest.raises(ValueError):
        T.create_uvw_tile_mapping(np.zeros((3, 3)), (0.0, 1.0, 1.0), freq)


def test_row_blocks_merge_like_the_reference_pool(T):
    """tiling_plan.py:114-147: binning row blocks with offsets and merging equals one pass."""
    rng = np.random.default_rng(9)
    uvw = rng.normal(size=(4001, 3)) * np.array([4000, 4000, 300.0])
    freq = 856e6 + np.arange(64) * 214e6 / 256
    ts = (500.0, 500.0, 1000.0)
    whole = T.create_uvw_tile_mapping(uvw, ts, freq)
    block = -(-len(uvw) // 3)
    parts = [T.create_uvw_tile_mapping_sequential(uvw[a:a + block], ts, freq, row_offset=a)
             for a in range(0, len(uvw), block)]
    merged = T.merge_tile_mappings(parts)
    assert list(merged) == list(whole) and merged == whole


def test_full_size_ska_low_binning_properties(T):
    """BASELINE.json configs[2] (130,816 baselines x 32 times x 128 channels = 5.36e8 samples): too big for
    the oracle, so check what must hold at any size -- every sample covered exactly once, slices
    sorted by (irow, chan_start) inside each tile, tile keys unique and in first-appearance order,
    and the tile index of sampled slices recomputed on the host with the reference formula."""
    from ska_sdp_cip import synthetic as S
    cfg = S.ska_low()
    uvw, freq = S.make_uvw(cfg), cfg.frequencies()
    ts = (200.0, 200.0, 400.0)
    dm = T.device_tile_mapping(uvw, ts, freq)
    coords, offsets, irow, c0, c1 = dm.arrays()
    assert offsets[0] == 0 and offsets[-1] == dm.n_slices == len(irow) and np.all(np.diff(offsets) > 0)
    assert len(np.unique(coords, axis=0)) == dm.n_tiles
    length = (c1 - c0).astype(np.int64)
    assert length.min() > 0 and int(length.sum()) == cfg.n_vis
    # per row: the slices tile [0, nchan) without gap or overlap (checksum of checksums)
    per_row = np.bincount(irow, weights=length, minlength=cfg.n_rows)
    assert np.all(per_row == cfg.n_chan)
    sq = np.bincount(irow, weights=(c1.astype(np.float64) ** 2 - c0.astype(np.float64) ** 2), minlength=cfg.n_rows)
    assert np.all(sq == cfg.n_chan ** 2)  # sum (b^2 - a^2) telescopes to nchan^2 only for a gap-free cover
    tile_of_slice = np.repeat(np.arange(dm.n_tiles), np.diff(offsets))
    key = irow * (cfg.n_chan + 1) + c0
    same_tile = tile_of_slice[1:] == tile_of_slice[:-1]
    assert np.all(key[1:][same_tile] > key[:-1][same_tile])
    # first-appearance order: the first slice of each tile is ordered by (irow, chan_start)
    first = key[offsets[:-1]]
    assert np.all(np.diff(first) > 0)
    # sampled slices: both end channels have the tile's index, the neighbours outside do not
    rng = np.random.default_rng(1)
    pick = rng.integers(0, dm.n_slices, 20000)
    idx = lambda r, ch: np.floor((freq[ch] / 299792458.0)[:, None] * (uvw[r] / np.array(ts)) + 0.5).astype(np.int64)
    want = coords[tile_of_slice[pick]]
    assert np.array_equal(idx(irow[pick], c0[pick]), want)
    assert np.array_equal(idx(irow[pick], c1[pick] - 1), want)
    inner = pick[c0[pick] > 0]
    assert np.all(np.any(idx(irow[inner], c0[inner] - 1) != coords[tile_of_slice[inner]], axis=1))


# ---- reorder_by_uvw_tile on the GPU path (reference tests/uvw_tiling/test_uvw_reordering.py) ----

def assert_all_visibilities_accounted_for(reader, paths):
    """The reference's acceptance check: the (u, v, w) in wavelengths of every sample appears exactly once."""
    from ska_sdp_cip.uvw_tiling import Tile
    c = 299792458.0
    freqs = reader.channel_frequencies()
    want = np.sort((reader.uvw().reshape(-1, 1, 3) / c * freqs.reshape(1, -1, 1)).reshape(-1, 3), axis=0)
    got, n = [], 0
    for p in paths:
        t = Tile.load_npz(p)
        for i, (a, b) in enumerate(zip(t.channel_start_indices, t.channel_stop_indices)):
            got.append(t.uvw[i] / c * freqs[a:b].reshape(-1, 1))
            n += b - a
    assert n == len(want)
    assert np.array_equal(want, np.sort(np.concatenate(got), axis=0))


@pytest.mark.gpu
def test_reorder_by_uvw_tile_matches_reference_golden(T, golden, tmp_path):
    from test_host_logic import check_reorder_against_golden, reorder_case
    from ska_sdp_cip.uvw_tiling import reorder_by_uvw_tile
    reader, tile_size, intervals, limit = reorder_case()
    paths = reorder_by_uvw_tile(reader, tile_size, tmp_path, None, num_time_intervals=intervals,
                                max_vis_per_chunk=limit)
    assert sorted(p.name for p in paths) == sorted(p.name for p in tmp_path.iterdir())
    check_reorder_against_golden(paths, golden("reorder_small"))
    assert_all_visibilities_accounted_for(reader, paths)


@pytest.mark.gpu
def test_reorder_with_a_stub_client_and_the_reference_test_parameters(T, tmp_path):
    """tile size, interval count and chunk limit of the reference's own reorder test, through the dask branch
    with an in-process stand-in for the client."""
    from test_host_logic import make_reader
    from ska_sdp_cip.uvw_tiling import reorder_by_uvw_tile

    class Future:
        def __init__(self, value): self.value = value
        def result(self): return self.value

    class StubClient:
        def __init__(self): self.resources = []
        def scheduler_info(self): return {"workers": {"a": {}, "b": {}}}
        def submit(self, fn, *args, resources=None, **kw):
            self.resources.append(resources)
            return Future(fn(*[a.result() if isinstance(a, Future) else a for a in args], **kw))

    reader = make_reader(nrow=3000, nchan=16, seed=5, dense=True)
    client = StubClient()
    paths = reorder_by_uvw_tile(reader, (3000.0, 3000.0, 6000.0), tmp_path, client, num_time_intervals=4,
                                max_vis_per_chunk=10_000)
    assert {"processing_slots": 1} in client.resources
    assert_all_visibilities_accounted_for(reader, paths)
    from ska_sdp_cip.uvw_tiling import Tile
    assert max(Tile.load_npz(p).num_visibilities for p in paths) <= 10_000


@pytest.mark.gpu
def test_reorder_app_runs_in_process(T, tmp_path):
    from test_host_logic import make_reader
    from ska_sdp_cip.apps import uvw_reorder_app
    reader = make_reader(nrow=800, nchan=8, seed=6, dense=True)
    out = tmp_path / "tiles"
    paths = uvw_reorder_app.run_program(["-t", "2000", "2000", "4000", "-o", str(out), "-n", "2", "-m", "5000", "x.ms"],
                                        reader_factory=lambda path: reader)
    assert out.is_dir() and paths
    assert_all_visibilities_accounted_for(reader, paths)


# ---- gridding straight from the reordered tiles -------------------------------------------------

@pytest.mark.gpu
def test_dirty_image_from_tiles_equals_the_image_from_the_rows(T, tmp_path):
    """reorder_by_uvw_tile -> tile chunk files -> invert_tiles: the tiles are gridded in their own (jagged) layout
    and give the image of the original rows (unit weights, no flags) and the DFT's."""
    from oracle import dft
    from test_host_logic import make_reader
    from ska_sdp_cip import invert_measurement_set
    from ska_sdp_cip.invert import invert_tiles
    from ska_sdp_cip.uvw_tiling import Tile, reorder_by_uvw_tile
    nrow, nchan, npix = 2500, 24, 160
    rng = np.random.default_rng(8)
    uvw = rng.normal(size=(nrow, 3)) * np.array([900.0, 900.0, 120.0])
    freqs = 1.2e9 + 2.5e6 * np.arange(nchan)
    vis_i = (rng.normal(size=(nrow, nchan)) + 1j * rng.normal(size=(nrow, nchan))).astype(np.complex64)
    vis4 = np.zeros((nrow, nchan, 4), dtype=np.complex64)
    vis4[..., 0] = vis4[..., 3] = vis_i
    from ska_sdp_cip.measurement_set import ArrayMeasurementSetReader
    reader = ArrayMeasurementSetReader(uvw, freqs, vis4, np.zeros(vis4.shape, dtype=bool),
                                       np.ones(vis4.shape, dtype=np.float32))
    paths = reorder_by_uvw_tile(reader, (700.0, 700.0, 1400.0), tmp_path, None, num_time_intervals=3,
                                max_vis_per_chunk=4000)
    assert len(paths) > 10
    px = 0.4 / (np.abs(uvw[:, :2]).max() * freqs.max() / 299792458.0)
    asec = float(np.degrees(np.arcsin(px)) * 3600.0)
    img_tiles = invert_tiles(paths, freqs, npix, asec)
    img_rows = invert_measurement_set(reader, npix, asec)
    assert img_tiles.shape == (npix, npix) and img_tiles.dtype == np.float32
    assert dft.rel_rms(img_tiles, img_rows) <= 2e-6
    px_used = float(np.sin(np.radians(asec / 3600.0)))
    ref = dft.dft_dirty(uvw, freqs, vis_i, None, npix, npix, px_used, px_used) / vis_i.size
    assert dft.rel_rms(img_tiles, ref) <= 1e-4
    # Tile objects work like paths; an empty list gives an empty image
    tiles = [Tile.load_npz(p) for p in paths[:5]]
    part = invert_tiles(tiles, freqs, npix, asec)
    assert np.isfinite(part).all() and np.abs(part).max() > 0
    assert not invert_tiles([], freqs, npix, asec).any()
