#!/usr/bin/env python
"""Development battery run on a GPU box: every stage against the oracle, many plans, with timings.
Not a test (tests/ holds those); prints one line per case so a single gpurun call tells a lot."""
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ska-sdp-continuum-imaging-pipeline_b200"))

from oracle import dft, tiling, wstack  # noqa: E402
from ska_sdp_cip import synthetic as S  # noqa: E402
from ska_sdp_cip import wgridder  # noqa: E402
from ska_sdp_cip.uvw_tiling import device_tile_mapping  # noqa: E402


def case(name, fn):
    t = time.time()
    try:
        out = fn()
        print(f"[ok ] {name}: {out}  ({time.time() - t:.2f}s)", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"[ERR] {name}: {type(e).__name__}: {e}", flush=True)
        traceback.print_exc()


def random_problem(nrow=2000, nchan=8, nx=128, fov_scale=1.0, wscale=60.0, seed=1, uvscale=800.0):
    rng = np.random.default_rng(seed)
    uvw = rng.normal(size=(nrow, 3)) * np.array([uvscale, uvscale, wscale])
    freq = 1.3e9 + 1e6 * np.arange(nchan)
    vis = (rng.normal(size=(nrow, nchan)) + 1j * rng.normal(size=(nrow, nchan))).astype(np.complex64)
    wgt = rng.uniform(0.5, 1.5, size=(nrow, nchan)).astype(np.float32)
    wgt[rng.random(wgt.shape) < 0.05] = 0
    px = fov_scale * 0.4 / (np.abs(uvw[:, :2]).max() * freq.max() / 299792458.0)
    return uvw, freq, vis, wgt, nx, px


def run(prob, eps=1e-4, fp64=False, sigma=None, W=None, slots=None, ref=None):
    uvw, freq, vis, wgt, nx, px = prob
    for k, v in (("CIP_SIGMA", sigma), ("CIP_SUPPORT", W), ("CIP_MAX_PLANE_SLOTS", slots)):
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = str(v)
    img, info = wgridder.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=eps, do_wstacking=True,
                                  double_precision_accumulation=fp64, return_info=True)
    for k in ("CIP_SIGMA", "CIP_SUPPORT", "CIP_MAX_PLANE_SLOTS"):
        os.environ.pop(k, None)
    if ref is None:
        ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    p, r = info["plan"], info["report"]
    return (f"rel_rms={dft.rel_rms(img, ref):.3e} sigma={p['sigma']} W={p['support']} nu={p['nu']} "
            f"planes={p['nplanes']} block={p['plane_block']} nblocks={r['nblocks']} "
            f"grid={r['ms_grid']:.2f}ms bin={r['ms_bin']:.2f}ms fft={r['ms_fft']:.2f}ms tot={r['ms_total']:.2f}ms")


def main():
    import torch
    print("device", torch.cuda.get_device_name(0), flush=True)
    prob = random_problem()
    ref = dft.dft_dirty(prob[0], prob[1], prob[2], prob[3], prob[4], prob[4], prob[5], prob[5])
    case("small default fp32", lambda: run(prob, ref=ref))
    case("small default fp64", lambda: run(prob, fp64=True, ref=ref))
    for sg, W in ((2.0, 6), (2.0, 4), (2.0, 8), (1.5, 7), (1.25, 8), (1.75, 5), (1.5, 10)):
        case(f"small sigma={sg} W={W} fp32", lambda sg=sg, W=W: run(prob, sigma=sg, W=W, ref=ref))
    case("small sigma=2 W=8 fp64", lambda: run(prob, fp64=True, sigma=2.0, W=8, ref=ref))
    case("small sigma=1.5 W=12 fp64", lambda: run(prob, fp64=True, sigma=1.5, W=12, ref=ref))
    case("small eps=1e-6 fp64", lambda: run(prob, eps=1e-6, fp64=True, ref=ref))

    # oracle w-stacking with the same plan: stage-level agreement
    def vs_wstack():
        uvw, freq, vis, wgt, nx, px = prob
        img, info = wgridder.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=True,
                                      return_info=True)
        p = info["plan"]
        o = wstack.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, 1e-4, sigma=p["sigma"], W=p["support"])
        return f"rel_rms vs oracle wstack (same plan) = {dft.rel_rms(img, o):.3e}"
    case("small vs oracle wstack", vs_wstack)

    wide = random_problem(nrow=3000, nchan=4, nx=192, fov_scale=1.0, wscale=400.0, seed=2, uvscale=30.0)
    wref = dft.dft_dirty(wide[0], wide[1], wide[2], wide[3], wide[4], wide[4], wide[5], wide[5])
    case("wide-field many planes fp32", lambda: run(wide, ref=wref))
    case("wide-field plane blocks (12 slots)", lambda: run(wide, slots=12, ref=wref))
    case("wide-field plane blocks (8 slots)", lambda: run(wide, slots=8, ref=wref))
    case("wide-field fp64 eps=1e-5", lambda: run(wide, eps=1e-5, fp64=True, ref=wref))
    case("wide-field fp64 blocks(9)", lambda: run(wide, eps=1e-5, fp64=True, slots=9, ref=wref))

    dense = random_problem(nrow=4000, nchan=16, nx=128, seed=3, uvscale=3.0, wscale=1.0)
    dense = (dense[0], dense[1], dense[2], dense[3], dense[4], 2e-4)
    case("dense centre (items split)", lambda: run(dense))

    def empty():
        uvw, freq, vis, wgt, nx, px = prob
        img = wgridder.ms2dirty(uvw, freq, vis, np.zeros_like(wgt), nx, nx, px, px, epsilon=1e-4, do_wstacking=True)
        return f"all-zero weights -> max|img|={np.abs(img).max()}"
    case("all flagged", empty)

    def nowst():
        uvw, freq, vis, wgt, nx, px = prob
        uvw0 = uvw.copy()
        uvw0[:, 2] = 0
        img = wgridder.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=False)
        r = dft.dft_dirty(uvw0, freq, vis, wgt, nx, nx, px, px)
        l = (np.arange(nx) - nx // 2) * px
        n = np.sqrt(1 - l[:, None] ** 2 - l[None, :] ** 2)
        return f"do_wstacking=False rel_rms vs w=0 DFT*n = {dft.rel_rms(img, r * n):.3e}"
    case("no wstacking", nowst)

    # tiling
    def tile_case(nrow, nchan, ts, seed, zero_every=50):
        rng = np.random.default_rng(seed)
        uvw = rng.normal(size=(nrow, 3)) * np.array([4000, 4000, 300.0])
        if zero_every:
            uvw[::zero_every] = 0
        freq = 856e6 + np.arange(nchan) * 214e6 / 256
        t = time.time()
        dm = device_tile_mapping(uvw, ts, freq, row_offset=5)
        arrs = dm.arrays()
        tg = time.time() - t
        t = time.time()
        o = tiling.tile_mapping_arrays(uvw, ts, freq, row_offset=5)
        to = time.time() - t
        same = all(np.array_equal(a, b) for a, b in zip(arrs, o))
        vis = (rng.normal(size=(nrow, nchan)) + 1j * rng.normal(size=(nrow, nchan))).astype(np.complex64)
        flat, voff, suvw = dm.gather(vis, uvw)
        irow = arrs[2] - 5
        ok_u = np.array_equal(suvw, uvw[irow])
        exp = np.concatenate([vis[r, a:b] for r, a, b in zip(irow[:2000], arrs[3][:2000], arrs[4][:2000])])
        ok_v = np.array_equal(flat[:len(exp)], exp)
        return f"tiles={dm.n_tiles} slices={dm.n_slices} identical={same} gather uvw={ok_u} vis={ok_v} gpu={tg:.3f}s oracle={to:.3f}s"
    case("tiling 3000x64", lambda: tile_case(3000, 64, (300.0, 300.0, 600.0), 0))
    case("tiling 20000x256 small tiles", lambda: tile_case(20000, 256, (50.0, 50.0, 100.0), 1))
    case("tiling 1x1", lambda: tile_case(1, 1, (3000.0, 3000.0, 6000.0), 2, zero_every=0))
    case("tiling 200000x64", lambda: tile_case(200000, 64, (3000.0, 3000.0, 6000.0), 3))

    # a slice of the headline configuration: throughput smoke
    def mid(nt, fp64=False):
        cfg = S.ska_mid(n_times=nt)
        uvw = S.make_uvw(cfg)
        freq = cfg.frequencies()
        px = S.pixel_size_lm(S.ska_mid(), S.make_uvw(S.ska_mid(), (0, 1)))  # same pixel as the full config
        dev = torch.device("cuda", 0)
        vis, wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
        d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
        out = []
        for it in range(3):
            img, info = wgridder.ms2dirty(d_uvw, d_freq, vis, wgt, cfg.num_pixels, cfg.num_pixels, px, px,
                                          epsilon=1e-4, do_wstacking=True, double_precision_accumulation=fp64,
                                          return_info=True)
            r, p = info["report"], info["plan"]
            out.append(f"tot={r['ms_total']:.1f} scan={r['ms_scan']:.1f} bin={r['ms_bin']:.1f} grid={r['ms_grid']:.1f} "
                       f"fft={r['ms_fft']:.1f} wscr={r['ms_wscreen']:.1f} corr={r['ms_correct']:.1f}")
        _, _, flux, pi, pj = S.source_list(cfg, px)
        img = img.cpu().numpy() / info["scan"]["sum_wgt"]
        peaks = " ".join(f"{img[i, j]:.3f}/{f:.3f}" for i, j, f in zip(pi, pj, flux))
        nvis = cfg.n_vis
        return (f"nvis={nvis/1e6:.1f}M sigma={p['sigma']} W={p['support']} nu={p['nu']} planes={p['nplanes']} | "
                + " | ".join(out) + f" | Mvis/s(last)={nvis / r['ms_total'] / 1e3:.0f} | peaks img/flux: {peaks}")
    case("ska-mid 16 times (10.5M vis) device-resident", lambda: mid(16))
    case("ska-mid 64 times (42M vis) device-resident", lambda: mid(64))
    case("ska-mid 16 times fp64", lambda: mid(16, fp64=True))
    for sg, W in ((2.0, 6), (1.25, 8), (1.75, 6)):
        def forced(sg=sg, W=W):
            os.environ["CIP_SIGMA"], os.environ["CIP_SUPPORT"] = str(sg), str(W)
            try:
                return mid(64)
            finally:
                os.environ.pop("CIP_SIGMA"); os.environ.pop("CIP_SUPPORT")
        case(f"ska-mid 64 times sigma={sg} W={W}", forced)


if __name__ == "__main__":
    main()
