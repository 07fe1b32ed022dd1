#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the REFERENCE's own Python code in the build container.

The reference package is imported from /root/reference/src with stub modules for its three absent
third-party imports (dask.distributed, ducc0.wgridder, casacore.tables) and a patched
importlib.metadata.version; none of the stubs is ever called.  Only functions that are pure
numpy are exercised:

  tiling_*.npz   create_uvw_tile_mapping_sequential, create_uvw_tile_mapping (processes=3)
                 (src/ska_sdp_cip/uvw_tiling/tiling_plan.py:29-134) and
                 Tile._from_jagged_visibilities_slice (src/ska_sdp_cip/uvw_tiling/tile.py:83-115)
  stokes_*.npz   StokesIGridderInput.from_measurement_set_reader / effective_weights /
                 integrate_weighted_images (src/ska_sdp_cip/invert.py:72-116, 200-209)
  partition.npz  balanced_chunk_bounds (src/ska_sdp_cip/measurement_set.py:361-391)
  tile_ops.npz   split_tile, concatenate_tiles, rechunk_tiles_on_disk (src/ska_sdp_cip/uvw_tiling/tile.py:127-265)
  reorder_small.npz  reorder_time_interval + rechunk_tile_chunk_group run interval by interval
                 (src/ska_sdp_cip/uvw_tiling/reorder.py:129-183): per tile, the visibilities in file order
  task_metrics.json  TaskMetrics(...).to_json of a synthetic task stream (src/ska_sdp_cip/task_metrics.py)

/root/reference does not exist on the GPU box, so the outputs are committed as fixtures.
Run:  python tools/make_golden.py
"""
import importlib.metadata
import sys
import types
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
OUT = ROOT / "tests" / "golden"
REF_SRC = Path("/root/reference/src")


def import_reference():
    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    def _never(*a, **k):
        raise RuntimeError("third-party stub called")

    stub("dask")
    stub("dask.distributed", Client=object, get_worker=_never, as_completed=_never, wait=_never, Future=object)
    sys.modules["dask"].distributed = sys.modules["dask.distributed"]
    stub("ducc0")
    stub("ducc0.wgridder", ms2dirty=_never)
    stub("casacore")
    stub("casacore.tables", table=object)
    orig = importlib.metadata.version
    importlib.metadata.version = lambda name: "0.0.0-reference" if name == "ska_sdp_cip" else orig(name)
    sys.path.insert(0, str(REF_SRC))
    import ska_sdp_cip  # noqa: F401  (the reference, not this repo's package)
    assert Path(ska_sdp_cip.__file__).is_relative_to(REF_SRC), ska_sdp_cip.__file__
    return ska_sdp_cip


def mapping_to_arrays(mapping):
    """dict order and list order preserved."""
    coords = np.array(list(mapping.keys()), dtype=np.int64).reshape(-1, 3)
    sizes = [len(v) for v in mapping.values()]
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    flat = [s for v in mapping.values() for s in v]
    irow = np.array([s.irow for s in flat], dtype=np.int64)
    c0 = np.array([s.chan_start for s in flat], dtype=np.int32)
    c1 = np.array([s.chan_stop for s in flat], dtype=np.int32)
    return coords, offsets, irow, c0, c1


TILING_CASES = {
    # name: (nrow, nchan, tile_size, uvw scale, freq0, dfreq, row_offset, seed, zero_every)
    "meerkat_l": (3000, 256, (3000.0, 3000.0, 6000.0), (4000.0, 4000.0, 300.0), 856e6, 214e6 / 256, 0, 11, 50),
    "small_tiles": (400, 64, (50.0, 50.0, 100.0), (900.0, 900.0, 80.0), 1.3e9, 1e6, 7, 12, 0),
    "one_channel": (500, 1, (30.0, 30.0, 60.0), (600.0, 600.0, 50.0), 150e6, 0.0, 0, 13, 25),
    "descending_freq": (300, 48, (120.0, 100.0, 200.0), (2500.0, 2500.0, 300.0), 1.0e9, -2e6, 3, 14, 0),
    "half_integers": (64, 32, (1.0, 1.0, 1.0), None, 299792458.0, 299792458.0, 0, 15, 0),
}


def tiling_inputs(case):
    nrow, nchan, ts, scale, f0, df, off, seed, zero_every = TILING_CASES[case]
    rng = np.random.default_rng(seed)
    if scale is None:
        # uvw/tile * freq/c lands exactly on k + 0.5 for many samples: the +0.5/floor rounding edge
        uvw = (rng.integers(-8, 9, size=(nrow, 3)) * 0.5).astype(np.float64)
    else:
        uvw = rng.normal(size=(nrow, 3)) * np.array(scale)
    if zero_every:
        uvw[::zero_every] = 0.0
    freq = f0 + df * np.arange(nchan, dtype=np.float64)
    return uvw, ts, freq, off, seed


def main():
    ref = import_reference()
    from ska_sdp_cip.invert import StokesIGridderInput, integrate_weighted_images
    from ska_sdp_cip.measurement_set import balanced_chunk_bounds
    from ska_sdp_cip.uvw_tiling.tile import Tile
    from ska_sdp_cip.uvw_tiling.tiling_plan import create_uvw_tile_mapping, create_uvw_tile_mapping_sequential

    OUT.mkdir(parents=True, exist_ok=True)
    for case in TILING_CASES:
        uvw, ts, freq, off, seed = tiling_inputs(case)
        mapping = create_uvw_tile_mapping_sequential(uvw, ts, freq, row_offset=off)
        coords, offsets, irow, c0, c1 = mapping_to_arrays(mapping)
        extra = {}
        if off == 0:
            par = create_uvw_tile_mapping(uvw, ts, freq, processes=3)
            pc = mapping_to_arrays(par)
            assert all(np.array_equal(a, b) for a, b in zip(pc, (coords, offsets, irow, c0, c1))), \
                "reference: parallel and sequential mappings differ"
        # tile gather for the first few tiles
        rng = np.random.default_rng(seed + 100)
        vis = (rng.normal(size=(len(uvw), len(freq))) + 1j * rng.normal(size=(len(uvw), len(freq)))).astype(np.complex64)
        keys = list(mapping.keys())[:6]
        for t, key in enumerate(keys):
            slices = [type(s)(s.irow - off, s.chan_start, s.chan_stop) for s in mapping[key]]
            tile = Tile._from_jagged_visibilities_slice(vis, uvw, key, slices)
            extra[f"tile{t}_uvw"] = tile.uvw
            extra[f"tile{t}_vis"] = tile.visibilities
            extra[f"tile{t}_start"] = tile.channel_start_indices
            extra[f"tile{t}_stop"] = tile.channel_stop_indices
        np.savez_compressed(OUT / f"tiling_{case}.npz", uvw=uvw, tile_size=np.array(ts), freq=freq,
                            row_offset=np.int64(off), vis_seed=np.int64(seed + 100), coords=coords, offsets=offsets,
                            irow=irow, chan_start=c0, chan_stop=c1, n_gather_tiles=np.int64(len(keys)), **extra)
        print(f"tiling_{case}: {len(coords)} tiles, {len(irow)} slices")

    # ---- Stokes I / effective weights / integrate ----
    class FakeReader:
        def __init__(self, seed, nrow, nchan):
            rng = np.random.default_rng(seed)
            self.v = (rng.normal(size=(nrow, nchan, 4)) + 1j * rng.normal(size=(nrow, nchan, 4))).astype(np.complex64)
            self.f = rng.random((nrow, nchan, 4)) < 0.1
            self.w = rng.uniform(0.0, 2.0, size=(nrow, nchan, 4)).astype(np.float32)
            self.w[rng.random((nrow, nchan, 4)) < 0.05] = 0.0
            self.u = rng.normal(size=(nrow, 3)) * 1000.0
            self.fr = 1.0e9 + 1.0e6 * np.arange(nchan)

        def visibilities(self): return self.v
        def flags(self): return self.f
        def weights(self): return self.w
        def uvw(self): return self.u
        def channel_frequencies(self): return self.fr

    rd = FakeReader(21, 257, 9)
    gi = StokesIGridderInput.from_measurement_set_reader(rd)
    effw = gi.effective_weights()
    rng = np.random.default_rng(22)
    imgs = [(rng.normal(size=(16, 16)).astype(np.float32), np.float32(rng.uniform(1, 5))) for _ in range(3)]
    integ = integrate_weighted_images(imgs)
    np.savez_compressed(OUT / "stokes_small.npz", vis4=rd.v, flags4=rd.f, weights4=rd.w, vis_i=gi.visibilities,
                        flags_i=gi.flags, weights_i=gi.weights, effw=effw, effw_sum=np.asarray(effw.sum()),
                        images=np.stack([i for i, _ in imgs]), image_weights=np.array([w for _, w in imgs]),
                        integrated=integ)
    print("stokes_small:", gi.visibilities.dtype, gi.flags.dtype, gi.weights.dtype, effw.dtype, integ.dtype)

    # ---- partition bounds ----
    cases = [(0, 74214, 1), (0, 74214, 2), (0, 74214, 5), (0, 74214, 7), (0, 4, 1), (0, 4, 3), (0, 4, 4),
             (100, 1000, 13), (5, 6, 1)]
    arr = {f"b_{a}_{b}_{k}": np.array(list(balanced_chunk_bounds(a, b, k)), dtype=np.int64) for a, b, k in cases}
    np.savez_compressed(OUT / "partition.npz", **arr)
    print("partition:", len(arr), "cases")
    golden_tile_ops()
    golden_reorder()
    golden_task_metrics()
    del ref


def random_tile(Tile, seed, nrows, nchan=40):
    rng = np.random.default_rng(seed)
    start = rng.integers(0, nchan // 2, size=nrows)
    stop = start + rng.integers(1, nchan // 2, size=nrows)
    nvis = int((stop - start).sum())
    vis = (rng.normal(size=nvis) + 1j * rng.normal(size=nvis)).astype(np.complex64)
    return Tile(coords=(1, -2, 3), uvw=rng.normal(size=(nrows, 3)), visibilities=vis,
                channel_start_indices=start.astype(int), channel_stop_indices=stop.astype(int))


TILE_OPS_LIMITS = (1, 7, 19, 20, 64, 333, 10**6)


def golden_tile_ops():
    import tempfile
    from ska_sdp_cip.uvw_tiling.tile import Tile, concatenate_tiles, rechunk_tiles_on_disk, split_tile
    out = {}
    tile = random_tile(Tile, 31, 97)
    for limit in TILE_OPS_LIMITS:
        chunks = split_tile(tile, limit)
        out[f"split_{limit}_rows"] = np.array([c.num_rows for c in chunks], dtype=np.int64)
        out[f"split_{limit}_nvis"] = np.array([c.num_visibilities for c in chunks], dtype=np.int64)
    parts = [random_tile(Tile, 40 + i, n) for i, n in enumerate((5, 1, 30, 12, 3, 50))]
    cat = concatenate_tiles(parts)
    out["cat_uvw"], out["cat_vis"] = cat.uvw, cat.visibilities
    out["cat_start"], out["cat_stop"] = cat.channel_start_indices, cat.channel_stop_indices
    for limit in (50, 200, 10**6):
        with tempfile.TemporaryDirectory() as d:
            d = Path(d)
            paths = []
            for i, t in enumerate(parts):
                paths.append(d / f"in{i}.npz")
                t.save_npz(paths[-1])
            written = rechunk_tiles_on_disk(paths, d, "tile_x", max_vis_per_chunk=limit)
            out[f"rechunk_{limit}_names"] = np.array([w.name for w in written])
            out[f"rechunk_{limit}_nvis"] = np.array([Tile.load_npz(w).num_visibilities for w in written], dtype=np.int64)
            out[f"rechunk_{limit}_rows"] = np.array([Tile.load_npz(w).num_rows for w in written], dtype=np.int64)
    np.savez_compressed(OUT / "tile_ops.npz", **out)
    print("tile_ops:", len(out), "arrays")


REORDER_CASE = dict(nrow=600, nchan=24, seed=51, tile_size=(1500.0, 1500.0, 3000.0), intervals=3, limit=700)


def reorder_reader(seed, nrow, nchan):
    rng = np.random.default_rng(seed)
    uvw = rng.normal(size=(nrow, 3)) * np.array([1500.0, 1500.0, 200.0])
    vis = (rng.normal(size=(nrow, nchan, 4)) + 1j * rng.normal(size=(nrow, nchan, 4))).astype(np.complex64)
    freqs = 1.0e9 + 2.0e6 * np.arange(nchan)
    return uvw, vis, freqs


def golden_reorder():
    """The reference's two reorder passes without dask: tile files per interval in time order, then the rechunk
    of each tile with its interval files taken in interval order (the reference globs them in directory order,
    which is not defined; the per-tile visibility SET is)."""
    import tempfile
    from ska_sdp_cip.measurement_set import balanced_chunk_bounds
    from ska_sdp_cip.uvw_tiling import reorder as R
    from ska_sdp_cip.uvw_tiling.tile import Tile, rechunk_tiles_on_disk
    from ska_sdp_cip.uvw_tiling.tiling_plan import create_uvw_tile_mapping_sequential
    c = REORDER_CASE
    uvw, vis, freqs = reorder_reader(c["seed"], c["nrow"], c["nchan"])

    class IntervalReader:
        def __init__(self, a, b): self.a, self.b = a, b
        def uvw(self): return uvw[self.a:self.b]
        def visibilities(self): return vis[self.a:self.b]

    out = {}
    with tempfile.TemporaryDirectory() as d:
        d = Path(d)
        present = []
        for i, (a, b) in enumerate(balanced_chunk_bounds(0, c["nrow"], c["intervals"])):
            rd = IntervalReader(a, b)
            mapping = create_uvw_tile_mapping_sequential(rd.uvw(), c["tile_size"], freqs)
            for key in R.reorder_time_interval(rd, mapping, d, interval_index=i):
                if key not in present:
                    present.append(key)
        out["coords"] = np.array(present, dtype=np.int64)
        for t, key in enumerate(present):
            u, v, w = key
            base = f"tile_iu{u:+03d}_iv{v:+03d}_iw{w:+03d}"
            inputs = sorted(d.glob(base + "_interval*.npz"))
            written = rechunk_tiles_on_disk(inputs, d, base, max_vis_per_chunk=c["limit"])
            tiles = [Tile.load_npz(p) for p in written]
            out[f"t{t}_names"] = np.array([p.name for p in written])
            out[f"t{t}_vis"] = np.concatenate([x.visibilities for x in tiles])
            out[f"t{t}_uvw"] = np.concatenate([x.uvw for x in tiles])
            out[f"t{t}_start"] = np.concatenate([x.channel_start_indices for x in tiles])
            out[f"t{t}_nvis"] = np.array([x.num_visibilities for x in tiles], dtype=np.int64)
    np.savez_compressed(OUT / "reorder_small.npz", **out)
    print("reorder_small:", len(present), "tiles")


TASK_STREAM = [
    {"key": "from_measurement_set_reader-0a1b", "worker": "tcp://10.0.0.1:4000", "status": "OK",
     "startstops": ({"action": "compute", "start": 10.0, "stop": 11.5},)},
    {"key": "worker_ducc_invert-77ff", "worker": "tcp://10.0.0.2:4000", "status": "OK",
     "startstops": ({"action": "transfer", "start": 11.5, "stop": 11.75, "source": "tcp://10.0.0.1:4000"},
                    {"action": "compute", "start": 12.0, "stop": 15.0})},
    {"key": "integrate_weighted_images-multi-dash-9c", "worker": "tcp://10.0.0.1:4000", "status": "error",
     "startstops": ({"action": "compute", "start": 15.0, "stop": 15.25},
                    {"action": "transfer", "start": 14.5, "stop": 14.75})},
]


def golden_task_metrics():
    import json
    from ska_sdp_cip.task_metrics import TaskMetrics
    (OUT / "task_metrics.json").write_text(json.dumps(
        {"stream": TASK_STREAM, "expected": json.loads(TaskMetrics(TASK_STREAM).to_json())}, indent=1))
    print("task_metrics: ok")


if __name__ == "__main__":
    main()
