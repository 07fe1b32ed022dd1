"""Host-side mirror of the reference interface (no GPU): partitioning, chunked reads, result types,
Tile containers, synthetic generators.  The cases follow the reference's own tests
(tests/test_measurement_set_partition_indices.py, tests/test_measurement_set_chunked_read.py)."""
from pathlib import Path

import numpy as np
import pytest

from oracle import partition as opart

from ska_sdp_cip import measurement_set as MS
from ska_sdp_cip import synthetic as S
from ska_sdp_cip.uvw_tiling import RowSliceId, Tile, merge_tile_mappings

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"


def make_reader(nrow=74214, nchan=4, seed=0, dense=False):
    rng = np.random.default_rng(seed)
    uvw = rng.normal(size=(nrow, 3)) * 1000.0
    freqs = 1.4e9 + 1e6 * np.arange(nchan)
    if dense:
        vis = (rng.normal(size=(nrow, nchan, 4)) + 1j * rng.normal(size=(nrow, nchan, 4))).astype(np.complex64)
        flags = rng.random((nrow, nchan, 4)) < 0.05
        wgt = rng.uniform(0.5, 1.5, (nrow, nchan, 4)).astype(np.float32)
    else:  # broadcast views: shape without memory
        vis = np.broadcast_to(np.complex64(1), (nrow, nchan, 4))
        flags = np.broadcast_to(np.False_, (nrow, nchan, 4))
        wgt = np.broadcast_to(np.float32(1), (nrow, nchan, 4))
    return MS.ArrayMeasurementSetReader(uvw, freqs, vis, flags, wgt)


# the reference's golden partition indices (tests/test_measurement_set_partition_indices.py:33-61)
REFERENCE_PARTITIONS = [
    (1, 1, [(0, 74214, 0, 4)]),
    (2, 3, [(0, 37107, 0, 2), (0, 37107, 2, 3), (0, 37107, 3, 4),
            (37107, 74214, 0, 2), (37107, 74214, 2, 3), (37107, 74214, 3, 4)]),
    (5, 1, [(0, 14843, 0, 4), (14843, 29686, 0, 4), (29686, 44529, 0, 4), (44529, 59372, 0, 4),
            (59372, 74214, 0, 4)]),
]


@pytest.mark.parametrize("row_chunks,freq_chunks,expected", REFERENCE_PARTITIONS)
def test_partition_indices_match_reference_golden(row_chunks, freq_chunks, expected):
    chunks = make_reader().partition(row_chunks, freq_chunks)
    assert [(c.row_start, c.row_end, c.channel_start, c.channel_end) for c in chunks] == expected
    assert opart.partition_bounds(74214, 4, row_chunks, freq_chunks) == expected


def test_partition_raises_on_excessive_num_chunks():
    r = make_reader()
    with pytest.raises(ValueError):
        r.partition(1_000_000, 1)
    with pytest.raises(ValueError):
        r.partition(1, 1_000_000)


def test_balanced_chunks_match_golden_fixture(golden):
    for key, want in golden("partition").items():
        _, a, b, k = key.split("_")
        assert np.array_equal(np.array(list(MS.balanced_chunk_bounds(int(a), int(b), int(k)))), want), key
    with pytest.raises(ValueError):
        list(MS.balanced_chunk_sizes(4, 5))
    with pytest.raises(ValueError):
        list(MS.balanced_chunk_sizes(0, 1))


@pytest.mark.parametrize("method", ["visibilities", "flags", "weights", "uvw", "channel_frequencies"])
@pytest.mark.parametrize("row_chunks,freq_chunks", [(1, 4), (2, 3), (7, 1)])
def test_chunked_read_reassembles_full_read(method, row_chunks, freq_chunks):
    # tests/test_measurement_set_chunked_read.py: concatenating chunk reads gives the full read
    r = make_reader(nrow=301, nchan=4, dense=True)
    full = getattr(r, method)()
    chunks = r.partition(row_chunks, freq_chunks)
    rows = []
    for i in range(row_chunks):
        parts = [getattr(c, method)() for c in chunks[i * freq_chunks:(i + 1) * freq_chunks]]
        if method == "uvw":
            rows.append(parts[0])
            assert all(np.array_equal(p, parts[0]) for p in parts)
        elif method == "channel_frequencies":
            rows.append(np.concatenate(parts))
        else:
            rows.append(np.concatenate(parts, axis=1))
    got = rows[0] if method == "channel_frequencies" else np.concatenate(rows, axis=0)
    assert np.array_equal(got, full)


def test_partition_of_a_chunk_stays_inside_it():
    r = make_reader(nrow=100, nchan=8)
    c = r.partition(3, 2)[3]
    sub = c.partition(2, 2)
    assert (sub[0].row_start, sub[-1].row_end) == (c.row_start, c.row_end)
    assert (sub[0].channel_start, sub[-1].channel_end) == (c.channel_start, c.channel_end)


def test_reader_validates_shapes():
    with pytest.raises(ValueError):
        MS.ArrayMeasurementSetReader(np.zeros((3, 3)), np.zeros(2), np.zeros((3, 2, 2), np.complex64),
                                     np.zeros((3, 2, 2), bool), np.zeros((3, 2, 2), np.float32))
    with pytest.raises(ValueError):
        MS.ArrayMeasurementSetReader(np.zeros((4, 3)), np.zeros(2), np.zeros((3, 2, 4), np.complex64),
                                     np.zeros((3, 2, 4), bool), np.zeros((3, 2, 4), np.float32))


def test_casacore_reader_errors(tmp_path):
    with pytest.raises(FileNotFoundError):
        MS.MeasurementSetReader("/nonexistent.ms")
    try:
        import casacore  # noqa: F401
        pytest.skip("casacore is installed")
    except ImportError:
        pass
    with pytest.raises(ImportError):
        MS.MeasurementSetReader(tmp_path)


def test_merge_tile_mappings_keeps_first_appearance_order():
    a = {(0, 0, 0): [RowSliceId(0, 0, 4)], (1, 0, 0): [RowSliceId(1, 0, 2)]}
    b = {(1, 0, 0): [RowSliceId(5, 2, 4)], (2, 2, 2): [RowSliceId(6, 0, 4)]}
    m = merge_tile_mappings([a, b])
    assert list(m) == [(0, 0, 0), (1, 0, 0), (2, 2, 2)]
    assert m[(1, 0, 0)] == [RowSliceId(1, 0, 2), RowSliceId(5, 2, 4)]
    assert a[(1, 0, 0)] == [RowSliceId(1, 0, 2)]  # inputs untouched


def test_tile_container_matches_reference_golden(golden, tmp_path):
    g = golden("tiling_small_tiles")
    from oracle import tiling
    off = int(g["row_offset"])
    uvw, freq = g["uvw"], g["freq"]
    rng = np.random.default_rng(int(g["vis_seed"]))
    vis = (rng.normal(size=(len(uvw), len(freq))) + 1j * rng.normal(size=(len(uvw), len(freq)))).astype(np.complex64)
    mapping = tiling.tile_mapping(uvw, tuple(g["tile_size"]), freq, row_offset=off)
    key = list(mapping)[0]
    slices = [RowSliceId(r - off, a, b) for r, a, b in mapping[key]]
    t = Tile._from_jagged_visibilities_slice(vis, uvw, key, slices)
    assert np.array_equal(t.uvw, g["tile0_uvw"]) and np.array_equal(t.visibilities, g["tile0_vis"])
    assert np.array_equal(t.channel_start_indices, g["tile0_start"])
    assert np.array_equal(t.channel_stop_indices, g["tile0_stop"])
    assert t.num_rows == len(slices) and t.num_visibilities == sum(b - a for _, a, b in slices)
    p = tmp_path / "tile.npz"
    t.save_npz(p)
    with np.load(p) as z:
        assert sorted(z.files) == ["channel_start_indices", "channel_stop_indices", "coords", "uvw", "visibilities"]
    u = Tile.load_npz(p)
    assert u.coords == key and np.array_equal(u.visibilities, t.visibilities) and np.array_equal(u.uvw, t.uvw)
    z = Tile._zeros((1, 2, 3), 4, 9)
    assert z.uvw.shape == (4, 3) and z.visibilities.dtype == np.complex64 and z.num_visibilities == 9


def test_synthetic_configurations_have_the_baseline_shapes():
    mid, low = S.ska_mid(), S.ska_low()
    assert (mid.n_baselines, mid.n_rows, mid.n_vis) == (10296, 2635776, 168689664)
    assert (low.n_baselines, low.n_rows, low.n_vis) == (130816, 4186112, 535822336)
    cfg4 = S.ska_mid(n_times=512, n_chan=256, num_pixels=16384, dfreq_hz=0.5e6)
    assert cfg4.n_vis == 1349517312
    uvw = S.make_uvw(S.ska_mid(n_times=4))
    assert uvw.shape == (4 * 10296, 3) and uvw.dtype == np.float64
    # a time range of a longer observation is the same rows as the slice of the whole
    whole = S.make_uvw(S.ska_mid(n_times=8))
    part = S.make_uvw(S.ska_mid(n_times=8), (4, 8))
    assert np.array_equal(whole[4 * 10296:], part)
    px = S.pixel_size_lm(mid, uvw)
    lam = 299792458.0 / mid.frequencies().max()
    assert np.abs(uvw[:, :2]).max() / lam * px <= 0.4 + 1e-12


def test_synthetic_visibilities_image_to_their_sources():
    """One source, no noise: the dirty image at the source pixel is the flux (the generator uses
    the sign convention of the gridder's measurement equation)."""
    cfg = S.meerkat_like(num_pixels=128)
    full = S.make_uvw(cfg)
    uvw = full[:: len(full) // 3000][:3000]
    freq = cfg.frequencies()
    px = S.pixel_size_lm(cfg, full)
    vis, wgt = S.make_vis_numpy(cfg, uvw, freq, px, noise=0.0, flag_fraction=0.0)
    assert vis.dtype == np.complex64 and wgt.dtype == np.float32 and vis.shape == (3000, len(freq))
    from oracle import dft
    l, m, flux, pi, pj = S.source_list(cfg, px)
    # rebuild the model from the documented formula and image each source separately
    finv = freq / 299792458.0
    model = np.zeros(vis.shape, dtype=np.complex128)
    for k in range(len(flux)):
        nm1 = -(l[k] ** 2 + m[k] ** 2) / (np.sqrt(1 - l[k] ** 2 - m[k] ** 2) + 1)
        ph = (uvw[:, 0] * l[k] + uvw[:, 1] * m[k] - uvw[:, 2] * nm1)[:, None] * finv[None, :]
        one = flux[k] * np.exp(-2j * np.pi * ph)
        model += one
        pix = np.array([pi[k] * cfg.num_pixels + pj[k]])
        peak = dft.dft_dirty(uvw, freq, one.astype(np.complex64), wgt, cfg.num_pixels, cfg.num_pixels, px, px,
                             pixels=pix)[0] / wgt.sum() * (1 + nm1)
        assert peak == pytest.approx(flux[k], rel=1e-5)
    assert np.allclose(vis, model.astype(np.complex64), atol=1e-5)


# ---- tile helpers, reorder passes, task metrics, command lines ------------------------

def _random_tile(seed, nrows, nchan=40):
    """Same generator as tools/make_golden.py::random_tile."""
    rng = np.random.default_rng(seed)
    start = rng.integers(0, nchan // 2, size=nrows)
    stop = start + rng.integers(1, nchan // 2, size=nrows)
    nvis = int((stop - start).sum())
    vis = (rng.normal(size=nvis) + 1j * rng.normal(size=nvis)).astype(np.complex64)
    return Tile((1, -2, 3), rng.normal(size=(nrows, 3)), vis, start.astype(int), stop.astype(int))


def test_split_and_concatenate_match_reference_golden(golden):
    from ska_sdp_cip.uvw_tiling import concatenate_tiles, split_tile
    g = golden("tile_ops")
    tile = _random_tile(31, 97)
    for limit in (1, 7, 19, 20, 64, 333, 10**6):
        chunks = split_tile(tile, limit)
        assert [c.num_rows for c in chunks] == g[f"split_{limit}_rows"].tolist(), limit
        assert [c.num_visibilities for c in chunks] == g[f"split_{limit}_nvis"].tolist(), limit
        back = concatenate_tiles(chunks)
        assert np.array_equal(back.visibilities, tile.visibilities) and np.array_equal(back.uvw, tile.uvw)
        assert all(c.coords == tile.coords for c in chunks)
    parts = [_random_tile(40 + i, n) for i, n in enumerate((5, 1, 30, 12, 3, 50))]
    cat = concatenate_tiles(parts)
    assert np.array_equal(cat.uvw, g["cat_uvw"]) and np.array_equal(cat.visibilities, g["cat_vis"])
    assert np.array_equal(cat.channel_start_indices, g["cat_start"])
    assert np.array_equal(cat.channel_stop_indices, g["cat_stop"])
    with pytest.raises(ValueError):
        concatenate_tiles([])
    other = _random_tile(1, 2)
    other.coords = (0, 0, 0)
    with pytest.raises(ValueError):
        concatenate_tiles([parts[0], other])
    empty = Tile._zeros((0, 0, 0), 0, 0)
    assert split_tile(empty, 10) == []


def test_rechunk_on_disk_matches_reference_golden(golden, tmp_path):
    from ska_sdp_cip.uvw_tiling import rechunk_tiles_on_disk
    g = golden("tile_ops")
    parts = [_random_tile(40 + i, n) for i, n in enumerate((5, 1, 30, 12, 3, 50))]
    for limit in (50, 200, 10**6):
        d = tmp_path / f"limit{limit}"
        d.mkdir()
        paths = []
        for i, t in enumerate(parts):
            paths.append(d / f"in{i}.npz")
            t.save_npz(paths[-1])
        written = rechunk_tiles_on_disk(paths, d, "tile_x", max_vis_per_chunk=limit)
        assert [w.name for w in written] == g[f"rechunk_{limit}_names"].tolist()
        tiles = [Tile.load_npz(w) for w in written]
        assert [t.num_visibilities for t in tiles] == g[f"rechunk_{limit}_nvis"].tolist()
        assert [t.num_rows for t in tiles] == g[f"rechunk_{limit}_rows"].tolist()
        assert np.array_equal(np.concatenate([t.visibilities for t in tiles]), g["cat_vis"])


def reorder_case():
    """Inputs of tests/golden/reorder_small.npz (tools/make_golden.py::reorder_reader)."""
    nrow, nchan = 600, 24
    rng = np.random.default_rng(51)
    uvw = rng.normal(size=(nrow, 3)) * np.array([1500.0, 1500.0, 200.0])
    vis = (rng.normal(size=(nrow, nchan, 4)) + 1j * rng.normal(size=(nrow, nchan, 4))).astype(np.complex64)
    freqs = 1.0e9 + 2.0e6 * np.arange(nchan)
    flags = np.zeros(vis.shape, dtype=bool)
    wgt = np.ones(vis.shape, dtype=np.float32)
    return MS.ArrayMeasurementSetReader(uvw, freqs, vis, flags, wgt), (1500.0, 1500.0, 3000.0), 3, 700


def check_reorder_against_golden(paths, g):
    by_tile = {}
    for p in sorted(paths):
        t = Tile.load_npz(p)
        by_tile.setdefault(t.coords, []).append((p.name, t))
    coords = [tuple(int(x) for x in c) for c in g["coords"]]
    assert set(by_tile) == set(coords)
    for i, key in enumerate(coords):
        names = [n for n, _ in by_tile[key]]
        tiles = [t for _, t in by_tile[key]]
        assert names == g[f"t{i}_names"].tolist(), key
        assert [t.num_visibilities for t in tiles] == g[f"t{i}_nvis"].tolist(), key
        assert np.array_equal(np.concatenate([t.visibilities for t in tiles]), g[f"t{i}_vis"]), key
        assert np.array_equal(np.concatenate([t.uvw for t in tiles]), g[f"t{i}_uvw"]), key
        assert np.array_equal(np.concatenate([t.channel_start_indices for t in tiles]), g[f"t{i}_start"]), key


def test_reorder_passes_with_host_mapping_match_reference_golden(golden, tmp_path):
    """The reference's calling convention (pre-computed TileMapping) on the host: interval files, then the
    rechunk of every tile -- bit-identical file contents and names."""
    from oracle import tiling
    from ska_sdp_cip.uvw_tiling import reorder as R
    reader, tile_size, intervals, limit = reorder_case()
    freqs = reader.channel_frequencies()
    present = set()
    for i, chunk in enumerate(reader.partition(intervals, 1)):
        mapping = {k: [RowSliceId(*s) for s in v] for k, v in tiling.tile_mapping(chunk.uvw(), tile_size, freqs).items()}
        present.update(R.reorder_time_interval(chunk, mapping, tmp_path, interval_index=i))
    assert (tmp_path / R._tile_filename(next(iter(present)), 0)).name.startswith("tile_iu")
    paths = [p for key in present for p in R.rechunk_tile_chunk_group(key, tmp_path, max_vis_per_chunk=limit)]
    assert not list(tmp_path.glob("*_interval*.npz"))
    check_reorder_against_golden(paths, golden("reorder_small"))


def test_task_metrics_match_reference_golden(tmp_path):
    import json
    from ska_sdp_cip.task_metrics import Task, TaskMetrics
    g = json.loads((GOLDEN_DIR / "task_metrics.json").read_text())
    stream = [dict(e, startstops=tuple(e["startstops"])) for e in g["stream"]]
    metrics = TaskMetrics(stream)
    assert len(metrics) == 3 and isinstance(metrics[0], Task)
    assert metrics[2].name == "integrate_weighted_images-multi-dash" and metrics[1].duration == 3.5
    metrics.save_json(tmp_path / "m.json", indent=4, sort_keys=True)
    assert json.loads((tmp_path / "m.json").read_text()) == g["expected"]
    staged = TaskMetrics.from_stage_reports([{"ms_scan": 1.0, "ms_bin": 2.0, "ms_total": 9.0, "launches": 5}], start=5.0)
    assert [t.name for t in staged] == ["scan", "bin"] and abs(staged[1].stop - 5.003) < 1e-12


def test_command_lines_parse_like_the_reference():
    from ska_sdp_cip.apps import pipeline_app, uvw_reorder_app
    a = pipeline_app.get_parser().parse_args(["-n", "2048", "-p", "5.0", "in.ms", "out.npy"])
    assert (a.num_pixels, a.pixel_size, a.dask_scheduler, a.row_chunks, a.freq_chunks) == (2048, 5.0, None, 1, None)
    a = pipeline_app.get_parser().parse_args(["--num-pixels", "64", "--pixel-size", "1", "-d", "tcp://x:1", "-rc", "2",
                                              "-fc", "3", "in.ms", "out"])
    assert (a.dask_scheduler, a.row_chunks, a.freq_chunks) == ("tcp://x:1", 2, 3)
    b = uvw_reorder_app.get_parser().parse_args(["-t", "3000", "3000", "6000", "-n", "4", "-m", "10000", "in.ms"])
    assert b.tile_size == [3000.0, 3000.0, 6000.0] and b.num_time_intervals == 4 and b.max_vis_per_chunk == 10000
    with pytest.raises(SystemExit):
        pipeline_app.get_parser().parse_args(["in.ms", "out.npy"])  # -n and -p are required


def test_w_slab_edges_partition_the_axis():
    from ska_sdp_cip.invert import w_slab_edges
    rng = np.random.default_rng(3)
    w = np.sort(rng.exponential(300.0, 100000))
    wmin, wmax = w[0], w[-1]
    counts, _ = np.histogram(w, bins=4096, range=(wmin, wmax))
    edges = w_slab_edges(counts, wmin, wmax, 8)
    assert len(edges) == 9 and edges[0] == 0.0 and np.isinf(edges[-1]) and all(a <= b for a, b in zip(edges, edges[1:]))
    pop = [np.count_nonzero((w >= a) & (w < b)) for a, b in zip(edges, edges[1:])]
    assert sum(pop) == len(w) and max(pop) < 1.1 * len(w) / 8 and min(pop) > 0.9 * len(w) / 8
    # a large plane cost equalises the |w| extent instead of the population
    wide = w_slab_edges(counts, wmin, wmax, 8, plane_cost=1e12, dw=1.0)
    ext = np.diff(np.clip(wide, wmin, wmax))
    assert ext.max() < 1.1 * (wmax - wmin) / 8
    assert w_slab_edges(counts, wmin, wmax, 1) == [0.0, float("inf")]


def test_measured_rebalancing_of_the_w_slabs_converges():
    """invert.rebalanced_bin_factor: ranks whose slab took longer than the mean get less next time.  Simulated ranks
    whose time is a fixed part (the W - 1 extra planes) plus a per-sample cost that the model does not know (2.5 times
    higher above a |w| threshold): the slowest rank starts 49 % above the mean and ends within 2 %."""
    from ska_sdp_cip.invert import rebalanced_bin_factor, w_slab_edges
    nb, wmin, wmax, nslabs = 4096, 0.0, 1000.0, 4
    counts = (1e6 * np.exp(-np.arange(nb) / 600.0)).astype(np.int64)
    true_cost = np.where(np.arange(nb) < 1500, 1.0, 2.5) * counts + 3e5
    fixed = 0.35 * true_cost.sum() / nslabs
    centres = wmin + (wmax - wmin) * (np.arange(nb) + 0.5) / nb

    def times(edges):
        slab = np.clip(np.searchsorted(np.array(edges[1:-1]), centres, side="right"), 0, nslabs - 1)
        return np.array([fixed + true_cost[slab == s].sum() for s in range(nslabs)])

    f = np.ones(nb)
    worst = []
    for _ in range(5):
        edges = w_slab_edges(counts, wmin, wmax, nslabs, plane_cost=1e8, dw=1.0, bin_factor=f)
        assert edges[0] == 0.0 and edges[-1] == float("inf") and all(a <= b for a, b in zip(edges, edges[1:]))
        t = times(edges)
        worst.append(t.max() / t.mean())
        f = rebalanced_bin_factor(f, edges, wmin, wmax, t)
        assert np.all(f > 0) and abs(f.mean() - 1.0) < 1e-12
    assert worst[0] > 1.4 and worst[-1] < 1.02 and all(b <= a + 1e-9 for a, b in zip(worst, worst[1:]))
    # balanced already (or a single slab): the factors stay as they are
    assert np.array_equal(rebalanced_bin_factor(f, edges, wmin, wmax, [1.0, 1.01, 0.99, 1.0]), f)
    assert np.array_equal(rebalanced_bin_factor(f, [0.0, float("inf")], wmin, wmax, [1.0]), f)
