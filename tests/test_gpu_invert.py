"""Parity of the CUDA invert path (through the C-ABI) with the oracle: the float64 DFT of the
measurement equation and the oracle's w-stacking restatement stage by stage.

Tolerances (BASELINE.json north_star): relative RMS <= 1e-4 for fp32 accumulation at
epsilon = 1e-4, <= 1e-5 for fp64 accumulation; tighter where epsilon is tighter."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import dft, stokes, wstack

from conftest import random_problem

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-4
TOL_FP64 = 1e-5


@pytest.fixture(scope="module")
def wg(built_lib):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from ska_sdp_cip import wgridder
    return wgridder


@pytest.fixture(scope="module")
def small():
    prob = random_problem()
    uvw, freq, vis, wgt, nx, px = prob
    return prob, dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)


@pytest.fixture(scope="module")
def wide():
    prob = random_problem(nrow=3000, nchan=4, nx=192, wscale=400.0, seed=2, uvscale=30.0)
    uvw, freq, vis, wgt, nx, px = prob
    return prob, dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)


class forced_plan:
    """Force (sigma, W) / the number of resident planes through the library's tuning variables."""

    def __init__(self, sigma=None, W=None, slots=None):
        self.env = {"CIP_SIGMA": sigma, "CIP_SUPPORT": W, "CIP_MAX_PLANE_SLOTS": slots}

    def __enter__(self):
        for k, v in self.env.items():
            if v is not None:
                os.environ[k] = str(v)

    def __exit__(self, *exc):
        for k in self.env:
            os.environ.pop(k, None)


def invert(wg, prob, **kw):
    uvw, freq, vis, wgt, nx, px = prob
    kw.setdefault("epsilon", 1e-4)
    return wg.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, do_wstacking=True, return_info=True, **kw)


def test_default_plan_fp32_meets_tolerance(wg, small):
    prob, ref = small
    img, info = invert(wg, prob)
    assert img.dtype == np.float32 and img.shape == ref.shape
    assert dft.rel_rms(img, ref) <= TOL_FP32
    assert info["report"]["launches"] > 0 and info["scan"]["nactive"] > 0


def test_default_plan_fp64_meets_tolerance(wg, small):
    prob, ref = small
    img, _ = invert(wg, prob, double_precision_accumulation=True)
    assert img.dtype == np.float64
    assert dft.rel_rms(img, ref) <= TOL_FP64


def test_complex128_input_selects_fp64(wg, small):
    (uvw, freq, vis, wgt, nx, px), ref = small
    img = wg.ms2dirty(uvw, freq, vis.astype(np.complex128), wgt.astype(np.float64), nx, nx, px, px, epsilon=1e-6,
                      do_wstacking=True)
    assert img.dtype == np.float64
    assert dft.rel_rms(img, ref) <= 1e-6


def test_complex128_input_is_read_natively(wg, small):
    """complex128 visibilities / float64 weights go through the boundary as they are (cip_invert_args.vis_c128):
    values that complex64 cannot hold (relative rounding 3e-8) must not cost accuracy, from host and from device
    memory, for float64 and float32 grids."""
    import torch
    (uvw, freq, vis, wgt, nx, px), _ = small
    rng = np.random.default_rng(11)
    vis128 = vis.astype(np.complex128) * (1.0 + 1e-3 * rng.standard_normal(vis.shape)) + 1e-5j * rng.standard_normal(vis.shape)
    wgt64 = wgt.astype(np.float64) * (1.0 + 1e-3 * rng.standard_normal(wgt.shape))
    ref = dft.dft_dirty(uvw, freq, vis128, wgt64, nx, nx, px, px)
    rounded = dft.dft_dirty(uvw, freq, vis128.astype(np.complex64), wgt64.astype(np.float32), nx, nx, px, px)
    floor32 = dft.rel_rms(rounded, ref)
    assert floor32 > 1e-8  # what a down-cast at the boundary would cost
    with forced_plan(2.0, 14):
        img = wg.ms2dirty(uvw, freq, vis128, wgt64, nx, nx, px, px, epsilon=1e-9, do_wstacking=True)
        dev = lambda a: torch.as_tensor(a).cuda()
        img_d = wg.ms2dirty(dev(uvw), dev(freq), dev(vis128), dev(wgt64), nx, nx, px, px, epsilon=1e-9, do_wstacking=True)
    assert img.dtype == np.float64 and img_d.dtype == torch.float64
    assert dft.rel_rms(img, ref) <= 2e-9 < floor32 / 5
    assert dft.rel_rms(img_d.cpu().numpy(), ref) <= 2e-9
    # masked samples and float32 grids with complex128 input (values are rounded once, into the records)
    mask = (rng.random(vis.shape) > 0.3).astype(np.uint8)
    ref_m = dft.dft_dirty(uvw, freq, vis128, wgt64 * mask, nx, nx, px, px)
    args = wg.make_args(uvw, freq, vis128, wgt64, mask, nx, nx, px, px, 1e-4, True, False, 0)
    assert args.vis_c128 == 1
    from ska_sdp_cip._cabi import Context, Plan, Report, ScanResult, _ptr, check, lib
    out = np.empty((nx, nx), dtype=np.float32)
    ctx = Context.get(-1)
    scan = ScanResult()
    with ctx.lock:
        check(lib().cip_invert(ctx.handle, C.byref(args), _ptr(out), C.byref(scan), C.byref(Plan()), C.byref(Report())))
    assert dft.rel_rms(out, ref_m) <= TOL_FP32
    assert scan.nactive == int(np.count_nonzero((wgt64 * mask) != 0))
    assert abs(scan.sum_wgt - float((wgt64 * mask).sum())) <= 1e-9 * abs(scan.sum_wgt)


# every branch of the gridder: the register-window kernels (W <= 8) at each support, the wide path
# (W > 8), each oversampling factor
@pytest.mark.parametrize("sigma,W,tol", [(2.0, 4, 2e-3), (2.0, 5, 3e-4), (2.0, 6, 1e-4), (2.0, 7, 1e-4), (2.0, 8, 1e-4),
                                         (1.5, 6, 3e-4), (1.5, 7, 1e-4), (1.5, 8, 1e-4), (1.75, 5, 5e-4),
                                         (1.75, 6, 1e-4), (1.25, 8, 2e-4), (1.5, 10, 1e-4), (2.0, 9, 1e-4)])
def test_forced_plans_fp32(wg, small, sigma, W, tol):
    prob, ref = small
    with forced_plan(sigma, W):
        img, info = invert(wg, prob)
    assert (info["plan"]["sigma"], info["plan"]["support"]) == (sigma, W)
    assert dft.rel_rms(img, ref) <= tol


@pytest.mark.parametrize("sigma,W,tol", [(2.0, 8, 1e-6), (1.5, 12, 1e-7), (2.0, 14, 1e-9), (1.25, 16, 1e-7)])
def test_forced_plans_fp64(wg, small, sigma, W, tol):
    prob, ref = small
    with forced_plan(sigma, W):
        img, _ = invert(wg, prob, double_precision_accumulation=True)
    assert dft.rel_rms(img, ref) <= tol


@pytest.mark.parametrize("eps,fp64", [(1e-2, False), (1e-3, False), (1e-4, False), (1e-5, True), (1e-6, True),
                                      (1e-8, True)])
def test_epsilon_is_met(wg, small, eps, fp64):
    prob, ref = small
    img, _ = invert(wg, prob, epsilon=eps, double_precision_accumulation=fp64)
    assert dft.rel_rms(img, ref) <= eps


def test_same_plan_agrees_with_oracle_wstacking(wg, small):
    """Stage-level check: with the same (sigma, W) the CUDA path and the oracle's w-stacking port differ
    only by float32 rounding, far below the kernel's aliasing error."""
    prob, _ = small
    uvw, freq, vis, wgt, nx, px = prob
    img, info = invert(wg, prob, double_precision_accumulation=True)
    p = info["plan"]
    o = wstack.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, 1e-4, fp64=True, sigma=p["sigma"], W=p["support"])
    assert dft.rel_rms(img, o) <= 1e-9


def test_wide_field_many_planes(wg, wide):
    prob, ref = wide
    img, info = invert(wg, prob)
    assert info["plan"]["nplanes"] > 20
    assert dft.rel_rms(img, ref) <= TOL_FP32
    img, _ = invert(wg, prob, epsilon=1e-5, double_precision_accumulation=True)
    assert dft.rel_rms(img, ref) <= TOL_FP64


@pytest.mark.parametrize("slots,fp64", [(12, False), (8, False), (9, True)])
def test_plane_blocks_give_the_same_image(wg, wide, slots, fp64):
    """When the plane stack does not fit, planes are processed in blocks; the image must not change."""
    prob, ref = wide
    with forced_plan(slots=slots):
        img, info = invert(wg, prob, double_precision_accumulation=fp64)
    assert info["report"]["nblocks"] > 1
    assert dft.rel_rms(img, ref) <= (TOL_FP64 if fp64 else TOL_FP32)


def test_dense_centre_splits_long_tiles(wg):
    prob = random_problem(nrow=4000, nchan=16, nx=128, seed=3, uvscale=3.0, wscale=1.0)
    prob = prob[:5] + (2e-4,)
    uvw, freq, vis, wgt, nx, px = prob
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    img, _ = invert(wg, prob)
    assert dft.rel_rms(img, ref) <= TOL_FP32


def test_rectangular_image_and_pixels(wg):
    uvw, freq, vis, wgt, _nx, px = random_problem(nrow=1500, nchan=4, seed=5)
    nx, ny, py = 96, 160, 0.7 * px
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, ny, px, py)
    img = wg.ms2dirty(uvw, freq, vis, wgt, nx, ny, px, py, epsilon=1e-4, do_wstacking=True)
    assert img.shape == (nx, ny)
    assert dft.rel_rms(img, ref) <= TOL_FP32


def test_negative_w_and_conjugate_symmetry(wg, small):
    """(u,v,w,V) and (-u,-v,-w,conj V) are the same measurement."""
    (uvw, freq, vis, wgt, nx, px), ref = small
    flip = np.arange(len(uvw)) % 2 == 0
    uvw2, vis2 = uvw.copy(), vis.copy()
    uvw2[flip] *= -1
    vis2[flip] = np.conj(vis2[flip])
    img = wg.ms2dirty(uvw2, freq, vis2, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=True)
    assert dft.rel_rms(img, ref) <= TOL_FP32


def test_no_weights_and_mask(wg, small):
    (uvw, freq, vis, wgt, nx, px), _ = small
    ref1 = dft.dft_dirty(uvw, freq, vis, np.ones_like(wgt), nx, nx, px, px)
    img = wg.ms2dirty(uvw, freq, vis, None, nx, nx, px, px, epsilon=1e-4, do_wstacking=True)
    assert dft.rel_rms(img, ref1) <= TOL_FP32
    mask = (np.random.default_rng(0).random(vis.shape) < 0.7).astype(np.uint8)
    ref2 = dft.dft_dirty(uvw, freq, vis, wgt * mask, nx, nx, px, px)
    img = wg.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=True, mask=mask)
    assert dft.rel_rms(img, ref2) <= TOL_FP32


def test_all_flagged_gives_zero_image(wg, small):
    (uvw, freq, vis, wgt, nx, px), _ = small
    img, info = wg.ms2dirty(uvw, freq, vis, np.zeros_like(wgt), nx, nx, px, px, epsilon=1e-4, do_wstacking=True,
                            return_info=True)
    assert info["scan"]["nactive"] == 0 and info["scan"]["sum_wgt"] == 0
    assert not img.any()


def test_single_visibility(wg):
    uvw = np.array([[120.0, -45.0, 30.0]])
    freq = np.array([1.0e9])
    vis = np.array([[1 + 2j]], dtype=np.complex64)
    nx, px = 64, 2e-4
    ref = dft.dft_dirty(uvw, freq, vis, None, nx, nx, px, px)
    img = wg.ms2dirty(uvw, freq, vis, None, nx, nx, px, px, epsilon=1e-4, do_wstacking=True)
    assert dft.rel_rms(img, ref) <= TOL_FP32


def test_autocorrelation_rows_at_the_origin(wg):
    """uvw = 0 rows (autocorrelations in the reference's test MS) land on the grid centre."""
    uvw, freq, vis, wgt, nx, px = random_problem(nrow=800, nchan=4, seed=8)
    uvw[::7] = 0.0
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    img = wg.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=True)
    assert dft.rel_rms(img, ref) <= TOL_FP32


def test_without_wstacking(wg, small):
    (uvw, freq, vis, wgt, nx, px), _ = small
    uvw0 = uvw.copy()
    uvw0[:, 2] = 0
    img = wg.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=False)
    l = (np.arange(nx) - nx // 2) * px
    n = np.sqrt(1 - l[:, None] ** 2 - l[None, :] ** 2)
    ref = dft.dft_dirty(uvw0, freq, vis, wgt, nx, nx, px, px) * n
    assert dft.rel_rms(img, ref) <= TOL_FP32


def test_linearity(wg, small):
    """The dirty image is linear in the visibilities (what chunked / sharded inversion relies on)."""
    (uvw, freq, vis, wgt, nx, px), _ = small
    rng = np.random.default_rng(11)
    vis_b = (rng.normal(size=vis.shape) + 1j * rng.normal(size=vis.shape)).astype(np.complex64)
    f = lambda v: wg.ms2dirty(uvw, freq, v, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=True,
                              double_precision_accumulation=True)
    a, b, ab = f(vis), f(vis_b), f((vis + 2 * vis_b).astype(np.complex64))
    assert dft.rel_rms(a + 2 * b, ab) <= 1e-6


def test_device_inputs_match_host_inputs(wg, small):
    import torch
    (uvw, freq, vis, wgt, nx, px), _ = small
    host = wg.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=True)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(a).to(dev)
    out = wg.ms2dirty(t(uvw), t(freq), t(vis), t(wgt), nx, nx, px, px, epsilon=1e-4, do_wstacking=True)
    assert out.is_cuda and out.dtype == torch.float32
    # same kernels on the same data; only the order of the reductions into the plane stack differs
    assert dft.rel_rms(out.cpu().numpy(), host) <= 2e-6


@pytest.mark.parametrize("chunks", [1, 3, 8])
def test_streamed_host_chunks_give_the_same_image(wg, small, chunks):
    """Host inputs are uploaded in row chunks that are binned and gridded as they arrive; the chunking
    must not change the image (beyond float32 summation order)."""
    (uvw, freq, vis, wgt, nx, px), ref = small
    os.environ["CIP_STREAM_CHUNKS"] = str(chunks)
    try:
        img, info = wg.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=True,
                                return_info=True)
    finally:
        os.environ.pop("CIP_STREAM_CHUNKS")
    assert info["report"]["nblocks"] == chunks
    assert dft.rel_rms(img, ref) <= TOL_FP32


def test_repeatable_to_rounding(wg, small):
    """Register windows are accumulated in a fixed order per work item; the order in which runs land in a bin
    (one cursor atomic per CTA of the fused runs pass) and the reductions of the windows into the global plane
    stack (red.global.add) are unordered, so two runs agree to float32 rounding in fp32 and to float64
    rounding in fp64."""
    prob, _ = small
    a, _ = invert(wg, prob)
    b, _ = invert(wg, prob)
    assert dft.rel_rms(a, b) <= 2e-6
    a, _ = invert(wg, prob, double_precision_accumulation=True)
    b, _ = invert(wg, prob, double_precision_accumulation=True)
    assert dft.rel_rms(a, b) <= 1e-14


def _with_env(name, value, fn):
    os.environ[name] = value
    try:
        return fn()
    finally:
        os.environ.pop(name)


@pytest.mark.parametrize("env", [("CIP_GRID_TC", "1"), ("CIP_GRID_MMA", "0"), ("CIP_GRID_MMA", "1"), ("CIP_DEVICE_CHUNKS", "3"), ("CIP_FFT_2D", "1"),
                                 ("CIP_FFT", "cufft"), ("CIP_BIN_TABLE_ENTRIES", "1")])
def test_kernel_variants_agree(wg, small, wide, env):
    """The alternatives kept behind environment switches compute the same sums in another order: the tcgen05
    gridder (accumulators in tensor memory, whose float32 accumulation truncates: looser agreement), FFMA2 rank-1
    updates and 3xTF32 mma.sync tiles instead of the two-term bfloat16 tiles this plan's kernel error allows (operands
    held to 2^-18: 2.8e-6 relative rms, profiles/r2z_accuracy_modes.txt), device inputs cut into chunks that are binned and gridded on two
    streams, cuFFT's 2-D plan and one w-screen pass / cuFFT 1-D transforms around a transposed crop instead of the own
    pruned transform with the fused w-screen, one binning pass per window (what a key table too large for one pass falls back to)."""
    name, value = env
    prob, ref = wide if name == "CIP_BIN_TABLE_ENTRIES" else small
    base, info = invert(wg, prob)
    if name == "CIP_DEVICE_CHUNKS":
        import torch
        dev = torch.device("cuda", 0)
        uvw, freq, vis, wgt, nx, px = prob
        t = lambda a: torch.from_numpy(a).to(dev)
        alt = _with_env(name, value, lambda: wg.ms2dirty(t(uvw), t(freq), t(vis), t(wgt), nx, nx, px, px, epsilon=1e-4,
                                                         do_wstacking=True)).cpu().numpy()
    else:
        alt, ainfo = _with_env(name, value, lambda: invert(wg, prob))
        if name == "CIP_BIN_TABLE_ENTRIES":
            assert ainfo["report"]["nblocks"] > info["report"]["nblocks"] >= 1
    assert dft.rel_rms(alt, base) <= {"CIP_GRID_TC": 6e-5, "CIP_GRID_MMA": 8e-6}.get(name, 2e-6)
    assert dft.rel_rms(alt, ref) <= TOL_FP32


def test_unreached_row_blocks_are_skipped(wg):
    """Planes of 128 MB and more count as 32 row blocks, and the blocks beyond the visibilities' |u| extent are neither
    cleared nor transformed along v nor read (CIP_ROW_BLOCKS=1 forces that on a small plane).  The skipped rows hold
    whatever the previous call left there -- here a problem that reaches every row -- and must not be read."""
    everywhere = random_problem(nrow=3000, nchan=8, nx=256, fov_scale=1.2, seed=5)
    centred = random_problem(nrow=3000, nchan=8, nx=256, fov_scale=0.4, seed=6)  # |u| px <= 0.16
    uvw, freq, vis, wgt, nx, px = centred
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    base, binfo = _with_env("CIP_ROW_BLOCKS", "0", lambda: invert(wg, centred))
    _with_env("CIP_ROW_BLOCKS", "1", lambda: invert(wg, everywhere))
    alt, ainfo = _with_env("CIP_ROW_BLOCKS", "1", lambda: invert(wg, centred))
    assert ainfo["plan"]["nu"] % 32 == 0 and ainfo["plan"] == binfo["plan"]
    # two transforms along v per plane (the first and the last rows) instead of one, and the extent kernel
    assert ainfo["report"]["launches"] - binfo["report"]["launches"] == ainfo["plan"]["nplanes"] + 1, "no row block was skipped"
    assert dft.rel_rms(alt, base) <= 2e-6
    assert dft.rel_rms(alt, ref) <= TOL_FP32
    full, finfo = _with_env("CIP_ROW_BLOCKS", "1", lambda: invert(wg, everywhere))
    uvw, freq, vis, wgt, nx, px = everywhere
    assert dft.rel_rms(full, dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)) <= TOL_FP32


def test_invalid_arguments_raise(wg, small):
    (uvw, freq, vis, wgt, nx, px), _ = small
    with pytest.raises(ValueError):
        wg.ms2dirty(uvw[:, :2], freq, vis, wgt, nx, nx, px, px)
    with pytest.raises(ValueError):
        wg.ms2dirty(uvw, freq, vis[:, :-1], wgt, nx, nx, px, px)
    with pytest.raises(ValueError):
        wg.ms2dirty(uvw, freq, vis, wgt[:-1], nx, nx, px, px)
    with pytest.raises(ValueError):
        wg.ms2dirty(uvw, freq, vis, wgt, nx + 1, nx, px, px, epsilon=1e-4)       # odd size
    with pytest.raises(ValueError):
        wg.ms2dirty(uvw, freq, vis, wgt, nx, nx, -px, px, epsilon=1e-4)          # negative pixel
    with pytest.raises(ValueError):
        wg.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=0.0)            # non-positive epsilon


# ---- the reference-facing surface (invert.py) ------------------------------------------------------

@pytest.fixture(scope="module")
def reader():
    from ska_sdp_cip import synthetic as S
    from ska_sdp_cip.measurement_set import ArrayMeasurementSetReader
    uvw, freq, vis, wgt, _nx, _px = random_problem(nrow=3000, nchan=8, seed=21, uvscale=300.0, wscale=30.0)
    vis4, flags4, wgt4 = S.four_pol(vis, wgt)
    return ArrayMeasurementSetReader(uvw, freq, vis4, flags4, wgt4)


def reference_invert(reader, npix, pix_asec):
    """The reference's invert_measurement_set arithmetic (invert.py:119-184) with the DFT as gridder."""
    v, f, w = stokes.stokes_i(reader.visibilities(), reader.flags(), reader.weights())
    effw = stokes.effective_weights(f, w)
    px = stokes.pixel_size_lm(pix_asec)
    img = dft.dft_dirty(reader.uvw(), reader.channel_frequencies(), v, effw, npix, npix, px, px)
    return img / effw.sum()


def test_stokes_prologue_is_bit_exact_with_reference_golden(built_lib, golden):
    from ska_sdp_cip.invert import StokesIGridderInput
    from ska_sdp_cip.measurement_set import ArrayMeasurementSetReader
    g = golden("stokes_small")
    nrow, nchan = g["vis4"].shape[:2]
    r = ArrayMeasurementSetReader(np.zeros((nrow, 3)), np.arange(nchan, dtype=float) + 1e9, g["vis4"], g["flags4"],
                                  g["weights4"])
    gi = StokesIGridderInput.from_measurement_set_reader(r)
    assert np.array_equal(gi.visibilities.cpu().numpy(), g["vis_i"])
    assert np.array_equal(gi.effective_weights().cpu().numpy(), g["effw"])


def test_invert_measurement_set(built_lib, reader):
    from ska_sdp_cip import invert_measurement_set
    npix, asec = 128, 20.0
    img = invert_measurement_set(reader, npix, asec)
    assert isinstance(img, np.ndarray) and img.shape == (npix, npix)  # the reference's own assertion
    assert img.dtype == np.float32
    assert dft.rel_rms(img, reference_invert(reader, npix, asec)) <= TOL_FP32


def test_pipeline_app_writes_the_dirty_image(built_lib, reader, tmp_path):
    """tests/test_pipeline_app.py:23-46 of the reference: run_program writes a (n, n) array next to the given
    path; here it must also equal invert_measurement_set's result."""
    from ska_sdp_cip import invert_measurement_set
    from ska_sdp_cip.apps.pipeline_app import run_program
    out = tmp_path / "image.npy"
    run_program(["-n", "128", "-p", "20.0", "x.ms", str(out)], reader_factory=lambda path: reader)
    data = np.load(out)
    assert data.shape == (128, 128)
    ref = invert_measurement_set(reader, 128, 20.0)
    assert dft.rel_rms(data, ref) <= 2e-6


@pytest.mark.parametrize("row_chunks,freq_chunks", [(1, None), (2, 3), (3, 1), (1, 8)])
def test_chunked_invert_is_chunk_invariant(built_lib, reader, row_chunks, freq_chunks):
    """tests/test_dask_invert_measurement_set.py:31-34: the chunked image equals the unchunked one
    (the reference asserts 1e-5 absolute on normalised images)."""
    from ska_sdp_cip import dask_invert_measurement_set, invert_measurement_set
    npix, asec = 128, 20.0
    whole = invert_measurement_set(reader, npix, asec)
    chunked = dask_invert_measurement_set(reader, None, npix, asec, row_chunks=row_chunks, freq_chunks=freq_chunks)
    assert chunked.shape == whole.shape and chunked.dtype == np.float32
    assert np.allclose(chunked, whole, atol=1e-5)
    assert dft.rel_rms(chunked, reference_invert(reader, npix, asec)) <= TOL_FP32


def test_dask_client_path_with_a_stub_client(built_lib, reader):
    """The dask code path (client.submit / futures / resources) driven by an in-process stand-in."""
    from ska_sdp_cip import dask_invert_measurement_set, invert_measurement_set

    class Future:
        def __init__(self, v):
            self.v = v

        def result(self):
            return self.v

    class Client:
        submitted = []

        def scheduler_info(self):
            return {"workers": {"a": {}, "b": {}}}

        def submit(self, fn, *args, resources=None, **kw):
            un = lambda a: [x.v if isinstance(x, Future) else x for x in a] if isinstance(a, list) else (
                a.v if isinstance(a, Future) else a)
            self.submitted.append((fn.__name__, resources))
            return Future(fn(*[un(a) for a in args], **kw))

    npix, asec = 128, 20.0
    client = Client()
    img = dask_invert_measurement_set(reader, client, npix, asec)
    names = [n for n, _ in client.submitted]
    assert names.count("worker_ducc_invert") == 2 and names[-1] == "integrate_weighted_images"
    assert all(r == {"processing_slots": 1} for n, r in client.submitted if n == "worker_ducc_invert")
    assert np.allclose(img, invert_measurement_set(reader, npix, asec), atol=1e-5)


def test_sharded_accumulate_then_correct_matches_single_call(built_lib, small):
    """cip_scan + cip_plan_make + cip_grid_accumulate (x shards) + cip_correct_image == cip_invert."""
    from ska_sdp_cip.invert import StokesIGridderInput, accumulate_shards, reduce_and_correct
    (uvw, freq, vis, wgt, nx, px), ref = small
    shards = [StokesIGridderInput(freq, None, uvw[a:b], vis[a:b], wgt[a:b]) for a, b in ((0, 700), (700, 2000))]
    image, plan, sumw = accumulate_shards(shards, nx, px)
    image = reduce_and_correct(image, plan, sumw, normalise=False)
    assert sumw == pytest.approx(float(wgt.sum(dtype=np.float64)), rel=1e-6)
    assert dft.rel_rms(image.cpu().numpy(), ref) <= TOL_FP32


def test_headline_configuration_slice_recovers_sources(built_lib):
    """A time slice of BASELINE.json configs[1] (full 4096^2 image): checked on sampled pixels against the
    DFT -- a size-independent property where the full DFT is unaffordable."""
    import torch
    from ska_sdp_cip import synthetic as S, wgridder
    cfg = S.ska_mid(n_times=2)
    uvw, freq = S.make_uvw(cfg), cfg.frequencies()
    full = S.ska_mid()
    px = S.pixel_size_lm(full, S.make_uvw(full, (0, 1)))
    dev = torch.device("cuda", 0)
    vis, wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
    img = wgridder.ms2dirty(torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev), vis, wgt, 4096, 4096, px,
                            px, epsilon=1e-4, do_wstacking=True).cpu().numpy()
    rng = np.random.default_rng(0)
    _, _, _, pi, pj = S.source_list(cfg, px)
    pix = np.concatenate([np.asarray(pi) * 4096 + np.asarray(pj), rng.integers(0, 4096 * 4096, 1016)])
    sel = slice(0, None, 16)  # every 16th row keeps the DFT to seconds; the image is linear, so grid the same rows
    vis_h, wgt_h = vis.cpu().numpy(), wgt.cpu().numpy()
    wsub = np.zeros_like(wgt_h)
    wsub[sel] = wgt_h[sel]
    img_sub = wgridder.ms2dirty(uvw, freq, vis_h, wsub, 4096, 4096, px, px, epsilon=1e-4, do_wstacking=True)
    ref = dft.dft_dirty(uvw[sel], freq, vis_h[sel], wgt_h[sel], 4096, 4096, px, px, pixels=pix)
    got = img_sub.reshape(-1)[pix]
    assert dft.rel_rms(got, ref) <= TOL_FP32
    assert np.isfinite(img).all() and np.abs(img).max() > 0


# ---- stage G1 on its own: the binning's records against the oracle's (oracle_wg_bin) -----------------------

def _product_records(wg, prob, fp64):
    """Decode cip_bin_inspect's output into absolute (i0, j0, p0, fu, fv, fw, re, im) per record."""
    import ctypes as C
    from ska_sdp_cip import _cabi
    uvw, freq, vis, wgt, nx, px = prob
    _, info = invert(wg, prob, double_precision_accumulation=fp64)
    p = info["plan"]
    plan = _cabi.Plan(**{k: p[k] for k, _ in _cabi.Plan._fields_})
    from ska_sdp_cip.wgridder import make_args
    args = make_args(uvw, freq, vis, wgt, None, nx, nx, px, px, 1e-4, True, fp64, 0)
    jobs, recs = _cabi.bin_inspect(_cabi.Context.get(-1), args, plan)
    W = p["support"]
    win_step = 8 - W + 1 if W <= 8 else 1
    ngroups = jobs["ngroups"] & 0x7fffffff
    assert np.all(ngroups >= 1) and np.all(ngroups <= 512)
    assert np.array_equal(jobs["group_begin"][1:], (jobs["group_begin"] + ngroups)[:-1])  # jobs tile the record array
    job_of = np.repeat(np.arange(len(jobs)), ngroups * 8)
    tu, tv, win = (jobs[k][job_of].astype(np.int64) for k in ("tu", "tv", "win"))
    if recs.shape[1] == 16:
        geo = recs[:, :8].copy().view("<u8").reshape(-1)
        val = recs[:, 8:].copy().view("<f4").reshape(-1, 2).astype(np.float64)
        i0r, j0r, p0r = geo & 7, (geo >> 3) & 7, (geo >> 6) & 15
        fu, fv, fw = (((geo >> s) & 0x3ffff) / 262144.0 for s in (10, 28, 46))
        snap = 9 - W  # every record of a job lies in the job's register window
        assert np.array_equal(i0r // snap, jobs["lu"][job_of]) and np.array_equal(j0r // snap, jobs["lv"][job_of])
    else:
        cell = recs[:, :4].copy().view("<u4").reshape(-1)
        i0r, j0r, p0r = cell & 31, (cell >> 5) & 31, (cell >> 10) & (15 if W <= 8 else 0x3fffff)
        if fp64:
            fu, fv, fw = (recs[:, o:o + 4].copy().view("<u4").reshape(-1) / 4294967296.0 for o in (4, 8, 12))
            val = recs[:, 16:32].copy().view("<f8").reshape(-1, 2)
        else:
            fu, fv, fw = (recs[:, o:o + 4].copy().view("<f4").reshape(-1).astype(np.float64) for o in (4, 8, 12))
            val = recs[:, 16:24].copy().view("<f4").reshape(-1, 2).astype(np.float64)
    i0 = tu * 8 - 8 + i0r.astype(np.int64)
    j0 = tv * 8 - 8 + j0r.astype(np.int64)
    p0 = win * win_step + p0r.astype(np.int64)
    keep = (val[:, 0] != 0) | (val[:, 1] != 0)   # padding and inactive samples carry zero values
    out = np.stack([i0, j0, p0, fu, fv, fw, val[:, 0], val[:, 1]], axis=1)[keep]
    return out, p


@pytest.mark.parametrize("which,fp64", [("small", False), ("wide", False), ("small", True)])
def test_binning_records_match_the_oracle(wg, small, wide, which, fp64):
    """Stage G1: every active sample appears exactly once, in the right (window, tile, lattice) job, with the oracle's
    support origin, first plane, fractional offsets (to the record format's resolution) and weighted,
    conjugated, n-shift-rotated value."""
    from oracle import plan as oplan
    prob, _ = small if which == "small" else wide
    uvw, freq, vis, wgt, nx, px = prob
    got, p = _product_records(wg, prob, fp64)
    opl = oplan.Plan(nx=nx, ny=nx, px=px, py=px, sigma=p["sigma"], W=p["support"], beta=p["beta"], nu=p["nu"], nv=p["nv"],
                     nshift=p["nshift"], dw=p["dw"], w0=p["w0"], nplanes=p["nplanes"], nm1_min=0.0)
    o = wstack.bin_records(uvw, freq, vis, wgt, opl)
    want = np.stack([o[k].astype(np.float64) for k in ("i0", "j0", "p0", "fu", "fv", "fw", "vr", "vi")], axis=1)
    assert len(got) == len(want) == int(np.count_nonzero((wgt != 0) & (vis != 0)))
    order = lambda a: a[np.lexsort((a[:, 7], a[:, 6], a[:, 2], a[:, 1], a[:, 0]))]
    got, want = order(got), order(want)
    # a fraction that rounds to 1.0 in the record is stored as the largest value below it; origins are exact
    assert np.array_equal(got[:, :3], want[:, :3])
    ftol = 2.0 ** -18 if (not fp64 and p["support"] <= 8) else (1e-9 if fp64 else 1e-6)
    assert np.abs(got[:, 3:6] - want[:, 3:6]).max() <= ftol
    vtol = 1e-12 if fp64 else 2e-7
    assert np.abs(got[:, 6:] - want[:, 6:]).max() <= vtol * np.abs(want[:, 6:]).max()


# ---- |w|-sharded invert (SURVEY 8e option 3) -------------------------------------------------------------------

def test_w_slabs_in_one_process_give_the_same_image(built_lib, wide):
    """accumulate_shards_by_w: global (sigma, W, dw, nshift), per-slab w0 / nplanes, images summed uncorrected."""
    from ska_sdp_cip.invert import StokesIGridderInput, accumulate_shards, accumulate_shards_by_w, reduce_and_correct
    (uvw, freq, vis, wgt, nx, px), ref = wide
    shards = [StokesIGridderInput(freq, None, uvw[a:b], vis[a:b], wgt[a:b]) for a, b in ((0, 1100), (1100, 3000))]
    img0, plan0, sumw0 = accumulate_shards(shards, nx, px)
    img0 = reduce_and_correct(img0, plan0, sumw0, normalise=False).cpu().numpy()
    for nslabs in (1, 3, 5):
        reports = []
        img, plan, sumw = accumulate_shards_by_w(shards, nx, px, nslabs=nslabs, reports=reports)
        assert sumw == pytest.approx(sumw0, rel=1e-12)
        img = reduce_and_correct(img, plan, sumw, normalise=False).cpu().numpy()
        # the slabs' planes sit at other w than the single plan's: the two images agree to the w kernel's error,
        # not to rounding
        assert dft.rel_rms(img, img0) <= (3e-6 if nslabs == 1 else 3e-5), nslabs
        assert dft.rel_rms(img, ref) <= TOL_FP32
        if nslabs > 1:  # every slab spans fewer planes than the whole
            assert max(r["nplanes"] for r in reports) < plan0.nplanes
            assert len({r["slab"] for r in reports}) == nslabs


def test_w_slab_argument_selects_samples(wg, wide):
    """cip_invert with (w_lo, w_hi): the images of complementary slabs add up to the full image, and the weight
    sums of the slabs to the full weight sum."""
    import ctypes as C
    from ska_sdp_cip import _cabi
    from ska_sdp_cip.wgridder import make_args
    (uvw, freq, vis, wgt, nx, px), ref = wide
    ctx = _cabi.Context.get(-1)
    aw = np.abs(uvw[:, 2:3]) * (freq[None, :] / 299792458.0)  # the library's own order of operations
    cut = float(np.median(aw[wgt != 0]))
    total, sums = np.zeros((nx, nx)), []
    for slab in ((0.0, cut), (cut, 1e300)):
        out = np.empty((nx, nx), dtype=np.float32)
        a = make_args(uvw, freq, vis, wgt, None, nx, nx, px, px, 1e-4, True, False, 0, w_slab=slab)
        scan = _cabi.ScanResult()
        with ctx.lock:
            _cabi.check(_cabi.lib().cip_invert(ctx.handle, C.byref(a), out.ctypes.data, C.byref(scan), None, None))
        inside = (aw >= slab[0]) & (aw < slab[1])
        assert scan.nactive == int(np.count_nonzero(inside & (wgt != 0)))
        sums.append(scan.sum_wgt)
        total += out
    assert sum(sums) == pytest.approx(float(wgt.sum(dtype=np.float64)), rel=1e-6)
    assert dft.rel_rms(total, ref) <= TOL_FP32
    counts = np.zeros(64, dtype=np.int64)
    a = make_args(uvw, freq, vis, wgt, None, nx, nx, px, px, 1e-4, True, False, 0)
    with ctx.lock:
        _cabi.check(_cabi.lib().cip_w_histogram(ctx.handle, C.byref(a), float(aw.min()), float(aw.max()), 64,
                                                counts.ctypes.data))
    want, _ = np.histogram(aw[wgt != 0], bins=64, range=(float(aw.min()), float(aw.max())))
    assert counts.sum() == want.sum() and np.abs(counts - want).sum() <= 4  # samples on a bin edge may fall either way


def test_reader_pipeline_blocks_give_the_same_stokes_input(built_lib, golden):
    """from_measurement_set_reader through the pinned double-buffered upload, for several block sizes (one block,
    ragged blocks, one row per block): bit-identical to the reference's Stokes conversion."""
    from ska_sdp_cip.invert import StokesIGridderInput
    from ska_sdp_cip.measurement_set import ArrayMeasurementSetReader
    g = golden("stokes_small")
    nrow, nchan = g["vis4"].shape[:2]
    rng = np.random.default_rng(4)
    uvw = rng.normal(size=(nrow, 3))
    r = ArrayMeasurementSetReader(uvw, np.arange(nchan, dtype=float) + 1e9, g["vis4"], g["flags4"], g["weights4"])
    for rows in (None, 100, 33, 1):
        gi = StokesIGridderInput.from_measurement_set_reader(r, upload_rows=rows)
        assert np.array_equal(gi.visibilities.cpu().numpy(), g["vis_i"]), rows
        assert np.array_equal(gi.effective_weights().cpu().numpy(), g["effw"]), rows
        assert np.array_equal(gi.uvw.cpu().numpy(), uvw)
    sub = r.partition(3, 2)[3]   # a chunk of the reference's partition reads only its rows and channels
    gi = StokesIGridderInput.from_measurement_set_reader(sub, upload_rows=20)
    rs, cs = slice(sub.row_start, sub.row_end), slice(sub.channel_start, sub.channel_end)
    assert np.array_equal(gi.visibilities.cpu().numpy(), g["vis_i"][rs, cs])
    assert np.array_equal(gi.effective_weights().cpu().numpy(), g["effw"][rs, cs])
