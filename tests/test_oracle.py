"""The CPU oracle against the golden vectors produced by the reference's own code
(tools/make_golden.py) and against itself (w-stacking restatement vs float64 DFT)."""
import numpy as np
import pytest

from oracle import dft, partition, plan as oplan, stokes, tiling, wstack

from conftest import random_problem

TILING_CASES = ["meerkat_l", "small_tiles", "one_channel", "descending_freq", "half_integers"]


@pytest.mark.parametrize("case", TILING_CASES)
def test_tiling_oracle_matches_reference_golden(golden, case):
    g = golden(f"tiling_{case}")
    got = tiling.tile_mapping_arrays(g["uvw"], tuple(g["tile_size"]), g["freq"], row_offset=int(g["row_offset"]))
    for name, a in zip(("coords", "offsets", "irow", "chan_start", "chan_stop"), got):
        assert np.array_equal(a, g[name]), name


@pytest.mark.parametrize("case", TILING_CASES)
def test_tiling_oracle_gather_matches_reference_golden(golden, case):
    g = golden(f"tiling_{case}")
    off = int(g["row_offset"])
    uvw, freq = g["uvw"], g["freq"]
    rng = np.random.default_rng(int(g["vis_seed"]))
    vis = (rng.normal(size=(len(uvw), len(freq))) + 1j * rng.normal(size=(len(uvw), len(freq)))).astype(np.complex64)
    mapping = tiling.tile_mapping(uvw, tuple(g["tile_size"]), freq, row_offset=off)
    for t, key in enumerate(list(mapping)[: int(g["n_gather_tiles"])]):
        slices = [(r - off, a, b) for r, a, b in mapping[key]]
        t_uvw, t_vis, t_a, t_b = tiling.gather_tile(vis, uvw, slices)
        assert np.array_equal(t_uvw, g[f"tile{t}_uvw"])
        assert np.array_equal(t_vis, g[f"tile{t}_vis"])
        assert np.array_equal(t_a, g[f"tile{t}_start"]) and np.array_equal(t_b, g[f"tile{t}_stop"])


def test_tiling_every_sample_covered_exactly_once():
    # the property the reference tests (tests/uvw_tiling/test_uvw_tiling_plan.py:7-32)
    rng = np.random.default_rng(0)
    uvw = rng.normal(size=(500, 3)) * np.array([4000, 4000, 300.0])
    freq = 856e6 + np.arange(256) * 214e6 / 256
    mapping = tiling.tile_mapping(uvw, (3000.0, 3000.0, 6000.0), freq)
    cover = np.zeros((500, 256), dtype=int)
    for slices in mapping.values():
        for r, a, b in slices:
            cover[r, a:b] += 1
    assert np.all(cover == 1)


def test_stokes_oracle_matches_reference_golden(golden):
    g = golden("stokes_small")
    v, f, w = stokes.stokes_i(g["vis4"], g["flags4"], g["weights4"])
    assert v.dtype == g["vis_i"].dtype and np.array_equal(v, g["vis_i"])
    assert np.array_equal(f, g["flags_i"])
    assert w.dtype == g["weights_i"].dtype and np.array_equal(w, g["weights_i"])
    e = stokes.effective_weights(f, w)
    assert np.array_equal(e, g["effw"]) and e.sum() == g["effw_sum"]
    integ = stokes.integrate_weighted_images(list(zip(g["images"], g["image_weights"])))
    assert np.array_equal(integ, g["integrated"])


def test_partition_oracle_matches_reference_golden(golden):
    g = golden("partition")
    for key, want in g.items():
        _, a, b, k = key.split("_")
        assert np.array_equal(np.array(partition.chunk_bounds(int(a), int(b), int(k))), want), key
    # the reference's own golden numbers (tests/test_measurement_set_partition_indices.py:51-61)
    assert partition.chunk_bounds(0, 74214, 5) == [(0, 14843), (14843, 29686), (29686, 44529), (44529, 59372),
                                                    (59372, 74214)]
    with pytest.raises(ValueError):
        partition.chunk_bounds(0, 4, 5)


@pytest.mark.parametrize("fp64,sigma,W,tol", [(True, 2.0, 8, 1e-6), (False, 1.5, 7, 5e-5), (True, 1.5, 10, 1e-6)])
def test_wstack_oracle_reproduces_dft(fp64, sigma, W, tol):
    uvw, freq, vis, wgt, nx, px = random_problem(nrow=600, nchan=4, nx=64)
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    img = wstack.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, 1e-4, fp64=fp64, sigma=sigma, W=W)
    assert dft.rel_rms(img, ref) < tol


def test_wstack_oracle_default_plan_meets_epsilon():
    uvw, freq, vis, wgt, nx, px = random_problem(nrow=800, nchan=4, nx=96, wscale=300.0, uvscale=40.0)
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    for eps in (1e-3, 1e-4):
        img = wstack.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, eps)
        assert dft.rel_rms(img, ref) < eps


def test_dft_oracle_point_source_and_sampled_pixels():
    # a unit point source at pixel (40, 20) gives D = sum of weights there (times 1/n)
    nx, px = 64, 1e-4
    rng = np.random.default_rng(3)
    uvw = rng.normal(size=(300, 3)) * np.array([500.0, 500.0, 50.0])
    freq = np.array([1.0e9, 1.1e9])
    l, m = (40 - 32) * px, (20 - 32) * px
    nm1 = -(l * l + m * m) / (np.sqrt(1 - l * l - m * m) + 1)
    ph = (uvw[:, 0] * l + uvw[:, 1] * m - uvw[:, 2] * nm1)[:, None] * freq[None, :] / 299792458.0
    vis = np.exp(-2j * np.pi * ph).astype(np.complex64)
    wgt = np.ones(vis.shape, dtype=np.float32)
    img = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    assert abs(img[40, 20] * (1 + nm1) - vis.size) < 1e-2
    assert img[40, 20] == img.max()
    pix = np.array([40 * nx + 20, 5, 777])
    sub = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px, pixels=pix)
    assert np.allclose(sub, img.reshape(-1)[pix], rtol=1e-12, atol=1e-9)


def test_plan_oracle_basic_properties():
    p = oplan.make_plan(1024, 1024, 1e-5, 1e-5, 1e-4, 0.0, 1000.0, 10 ** 6)
    assert p.nu >= 1024 * p.sigma - 1 and p.nu % 2 == 0
    assert oplan.good_size(p.nu) == p.nu
    assert p.nplanes == int(np.floor(1000.0 / p.dw + p.W))
    assert oplan.ES_TABLE[p.sigma][p.W][1] <= 1e-4 / 3
    # planes are centred on the data range
    assert abs((p.w0 + 0.5 * (p.nplanes - 1) * p.dw) - 500.0) < 1e-9


def test_torch_dft_checker_is_pinned_to_the_dft_oracle():
    """oracle/dft_torch.py (used on the device for the BASELINE-size parity checks) against oracle/dft.py."""
    from oracle import dft_torch
    nx, ny, px, py = 96, 64, 2.1e-4, 1.7e-4
    rng = np.random.default_rng(11)
    uvw = rng.normal(size=(700, 3)) * np.array([900.0, 700.0, 150.0])
    freq = 1.0e9 + 3.0e6 * np.arange(5)
    vis = (rng.normal(size=(700, 5)) + 1j * rng.normal(size=(700, 5))).astype(np.complex64)
    wgt = rng.uniform(0.0, 2.0, size=(700, 5)).astype(np.float32)
    pix = dft_torch.sample_pixels(nx, ny, 200, seed=3, always=[0, nx * ny - 1, (nx // 2) * ny + ny // 2])
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, ny, px, py, pixels=pix)
    got = dft_torch.dft_pixels(uvw, freq, vis, wgt, nx, ny, px, py, pix, rows_per_block=97)
    assert dft.rel_rms(got, ref) < 1e-12
    sub = dft_torch.dft_pixels(uvw, freq, vis, wgt, nx, ny, px, py, pix, row_step=7)
    assert dft.rel_rms(sub, dft.dft_dirty(uvw[::7], freq, vis[::7], wgt[::7], nx, ny, px, py, pixels=pix)) < 1e-12
    assert dft.rel_rms(dft_torch.dft_pixels(uvw, freq, vis, None, nx, ny, px, py, pix),
                       dft.dft_dirty(uvw, freq, vis, None, nx, ny, px, py, pixels=pix)) < 1e-12


def test_fast_cpu_baseline_reproduces_the_dft():
    """oracle/csrc/wgrid_fast.c (the timed CPU arm of bench.py) computes the same image as the float64 checker to
    float32 accuracy, also when the plane stack is processed in blocks."""
    from conftest import random_problem
    uvw, freq, vis, wgt, nx, px = random_problem(nrow=3000, nchan=4, nx=192, wscale=400.0, seed=2, uvscale=30.0)
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    fast = wstack.ms2dirty_fast(uvw, freq, vis, wgt, nx, nx, px, px, 1e-4)
    blocks = wstack.ms2dirty_fast(uvw, freq, vis, wgt, nx, nx, px, px, 1e-4, max_stack_bytes=3 * 8 * 288 * 288)
    assert dft.rel_rms(fast, ref) < 1e-4 and dft.rel_rms(blocks, ref) < 1e-4
    assert dft.rel_rms(fast, wstack.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, 1e-4)) < 5e-6
    # all records in the tiles around the uv origin: the path on which every thread works on a private copy of a dense tile
    uvw, freq, vis, wgt, nx, _ = random_problem(nrow=4000, nchan=16, nx=128, seed=3, uvscale=3.0, wscale=1.0)
    px = 2e-4
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    fast = wstack.ms2dirty_fast(uvw, freq, vis, wgt, nx, nx, px, px, 1e-4)
    assert dft.rel_rms(fast, ref) < 1e-4
    assert dft.rel_rms(fast, wstack.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, 1e-4)) < 3e-5  # float32 sums of 64,000 records
