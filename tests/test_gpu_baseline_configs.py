"""Parity at the sizes BASELINE.json names: the CUDA invert through the C-ABI on the FULL inputs of configs[1],
configs[2] and the per-GPU share of configs[3], checked on the 8 injected source pixels + 1024 random pixels
against the float64 DFT (oracle/dft_torch.py, evaluated on the device; pinned to oracle/dft.py by
tests/test_oracle.py).  Tolerance as in BASELINE.json north_star: relative RMS <= 1e-4 for float32 grids at
epsilon = 1e-4, <= 1e-5 for float64 accumulation.  PARITY UNPINNED against ducc0 (absent here)."""
import numpy as np
import pytest

from oracle import dft, dft_torch

pytestmark = pytest.mark.gpu

TOL_FP32, TOL_FP64 = 1e-4, 1e-5


@pytest.fixture(scope="module")
def dev(built_lib):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    return torch.device("cuda", 0)


def run_config(cfg, dev, *, chan=None, fp64=False, row_step=1):
    """Invert the synthetic configuration (channels ``chan`` of it) from device-resident inputs; returns
    (sampled image values, DFT values, info)."""
    import torch
    from ska_sdp_cip import synthetic as S, wgridder
    uvw = S.make_uvw(cfg)
    px = S.pixel_size_lm(cfg, uvw)
    freq = cfg.frequencies()
    if chan is not None:
        freq = freq[chan]
    vis, wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
    d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
    n = cfg.num_pixels
    img, info = wgridder.ms2dirty(d_uvw, d_freq, vis, wgt, n, n, px, px, epsilon=cfg.epsilon, do_wstacking=True,
                                  double_precision_accumulation=fp64, return_info=True)
    _, _, _, pi, pj = S.source_list(cfg, px)
    pix = dft_torch.sample_pixels(n, n, 1024, seed=0, always=np.asarray(pi) * n + np.asarray(pj))
    got = img.reshape(-1)[torch.from_numpy(pix).to(dev)].cpu().numpy()
    assert bool(torch.isfinite(img).all())
    del img
    torch.cuda.empty_cache()
    ref = dft_torch.dft_pixels(d_uvw, d_freq, vis, wgt, n, n, px, px, pix, device=dev, row_step=row_step)
    return got, ref, info


def test_configs1_full_input_matches_dft_on_sampled_pixels(dev):
    """BASELINE.json configs[1], all 256 time steps -- the bench's input, and the bench's plan."""
    from ska_sdp_cip import synthetic as S
    cfg = S.ska_mid()
    got, ref, info = run_config(cfg, dev)
    plan = info["plan"]
    assert (plan["sigma"], plan["support"], plan["nplanes"], plan["nu"], plan["nv"]) == (1.5, 7, 7, 6144, 6144)
    assert info["scan"]["nactive"] > 0.97 * cfg.n_vis
    assert dft.rel_rms(got, ref) <= TOL_FP32
    # the 8 injected sources stand out of the noise at their pixels
    assert np.all(got[:8] > 5 * np.std(got[8:]))


def test_configs2_ska_low_full_input_matches_dft_on_sampled_pixels(dev):
    """BASELINE.json configs[2]: 130,816 baselines x 32 times x 128 channels = 536 M visibilities, 8192^2."""
    from ska_sdp_cip import synthetic as S
    cfg = S.ska_low()
    assert cfg.n_vis == 535_822_336
    got, ref, info = run_config(cfg, dev)
    assert info["plan"]["nu"] >= 8192 * 1.25 and info["plan"]["nplanes"] >= 7
    assert dft.rel_rms(got, ref) <= TOL_FP32


def test_configs3_per_gpu_share_matches_dft_on_sampled_pixels(dev):
    """BASELINE.json configs[3], the share of one of 8 GPUs under the reference's partition(1, freq_chunks=8)
    (32 of the 256 channels of all 5.27 M rows = 169 M visibilities), 16384^2 image: 24576^2 planes."""
    from ska_sdp_cip import synthetic as S
    from ska_sdp_cip.measurement_set import balanced_chunk_bounds
    cfg = S.ska_mid(n_times=512, n_chan=256, num_pixels=16384, dfreq_hz=0.5e6)
    assert cfg.n_vis == 1_349_517_312
    c0, c1 = list(balanced_chunk_bounds(0, cfg.n_chan, 8))[3]
    got, ref, info = run_config(cfg, dev, chan=slice(c0, c1))
    assert info["plan"]["nu"] >= 16384 * 1.25 and info["plan"]["nplanes"] >= 10
    assert dft.rel_rms(got, ref) <= TOL_FP32


def test_configs4_w_slab_share_fp64_matches_dft_on_sampled_pixels(dev):
    """BASELINE.json configs[4]: high-w SKA-Low, 32768^2 image, epsilon 1e-5, float64 grids, w planes sharded over 8
    GPUs.  One time step of the configuration (130,816 rows x 128 channels = 16.7 M samples); of its 8 |w| slabs of
    equal cost (invert.w_slab_edges) the share of ONE GPU is gridded -- tens of planes of 40960^2 complex128
    processed in blocks of the few that fit 180 GB -- and checked against the DFT of the slab's samples."""
    import torch
    from ska_sdp_cip import _cabi, synthetic as S, wgridder
    from ska_sdp_cip.invert import PLANE_COST_PER_CELL, w_slab_edges
    cfg = S.ska_low(high_w=True, num_pixels=32768)
    assert cfg.fp64 and cfg.epsilon == 1e-5
    uvw_all = S.make_uvw(cfg)
    px = S.pixel_size_lm(cfg, uvw_all)
    t = cfg.n_times // 2
    uvw = np.ascontiguousarray(uvw_all[t * cfg.n_baselines:(t + 1) * cfg.n_baselines])
    freq = cfg.frequencies()
    vis, wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
    aw = np.abs(uvw[:, 2:3]) * (freq[None, :] / 299792458.0)
    wmin, wmax = float(aw.min()), float(aw.max())
    n = cfg.num_pixels
    common = _cabi.plan_make(n, n, px, px, cfg.epsilon, wmin, wmax, aw.size, fp64=True)
    counts, _ = np.histogram(aw, bins=4096, range=(wmin, wmax))
    edges = w_slab_edges(counts, wmin, wmax, 8, plane_cost=PLANE_COST_PER_CELL * common.nu * common.nv, dw=common.dw)
    slab = (edges[3], edges[4])
    d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
    img, info = wgridder.ms2dirty(d_uvw, d_freq, vis, wgt, n, n, px, px, epsilon=cfg.epsilon, do_wstacking=True,
                                  double_precision_accumulation=True, return_info=True, w_slab=slab)
    plan, rep = info["plan"], info["report"]
    assert plan["accumulate_fp64"] == 1 and plan["nu"] >= 40960
    assert plan["nplanes"] < common.nplanes / 4            # the slab spans a fraction of the 600+ planes of the whole
    assert rep["nblocks"] >= 2 and plan["plane_block"] < plan["nplanes"]   # processed in plane blocks
    _, _, _, pi, pj = S.source_list(cfg, px)
    pix = dft_torch.sample_pixels(n, n, 1024, seed=0, always=np.asarray(pi) * n + np.asarray(pj))
    got = img.reshape(-1)[torch.from_numpy(pix).to(dev)].cpu().numpy()
    del img
    torch.cuda.empty_cache()
    inside = torch.from_numpy((aw >= slab[0]) & (aw < slab[1])).to(dev)
    assert int(inside.sum()) == info["scan"]["nactive"] + int((inside & (wgt == 0)).sum())
    ref = dft_torch.dft_pixels(d_uvw, d_freq, vis, wgt * inside, n, n, px, px, pix, device=dev)
    assert dft.rel_rms(got, ref) <= TOL_FP64
