"""Run one BASELINE.json configuration (per-GPU share) through ms2dirty on device inputs and print the plan,
the stage times and the memory high-water mark:  python tools/run_config.py cfg4|cfg3|cfg5|cfg5slab [n_times] [npix]
(cfg5slab: configs[4] at 32768^2 -- one time step, one of 8 cost-balanced |w| slabs, float64, in plane blocks)"""
import json, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "ska-sdp-continuum-imaging-pipeline_b200"))
import numpy as np, torch
from ska_sdp_cip import synthetic as S, wgridder

which = sys.argv[1]
if which == "cfg4":    # 1/8 of configs[3]: 64 of 512 times, 256 channels, 16384^2
    cfg = S.ska_mid(n_times=int(sys.argv[2]) if len(sys.argv) > 2 else 64, n_chan=256, num_pixels=16384, dfreq_hz=0.5e6)
    full = S.ska_mid(n_times=512, n_chan=256, num_pixels=16384, dfreq_hz=0.5e6)
elif which == "cfg3":  # configs[2]: SKA-Low 8192^2
    cfg = S.ska_low(n_times=int(sys.argv[2]) if len(sys.argv) > 2 else 32)
    full = cfg
elif which == "cfg5slab":
    cfg = S.ska_low(high_w=True, num_pixels=32768)
    full = cfg
else:                  # configs[4] share: high-w SKA-Low, fp64, eps 1e-5
    cfg = S.ska_low(n_times=int(sys.argv[2]) if len(sys.argv) > 2 else 4, num_pixels=int(sys.argv[3]) if len(sys.argv) > 3 else 8192, high_w=True)
    full = cfg
dev = torch.device("cuda", 0)
uvw_full = S.make_uvw(full, (0, full.n_times))
px = S.pixel_size_lm(full, uvw_full)
uvw = S.make_uvw(full, (0, cfg.n_times))
freq = cfg.frequencies()
slab = None
if which == "cfg5slab":
    from ska_sdp_cip import _cabi
    from ska_sdp_cip.invert import PLANE_COST_PER_CELL, w_slab_edges
    t = cfg.n_times // 2
    uvw = np.ascontiguousarray(uvw_full[t * cfg.n_baselines:(t + 1) * cfg.n_baselines])
    aw = np.abs(uvw[:, 2:3]) * (freq[None, :] / 299792458.0)
    common = _cabi.plan_make(cfg.num_pixels, cfg.num_pixels, px, px, cfg.epsilon, float(aw.min()), float(aw.max()), aw.size, fp64=True)
    counts, _ = np.histogram(aw, bins=4096, range=(float(aw.min()), float(aw.max())))
    edges = w_slab_edges(counts, float(aw.min()), float(aw.max()), 8, plane_cost=PLANE_COST_PER_CELL * common.nu * common.nv, dw=common.dw)
    slab = (edges[3], edges[4])
    print("whole configuration:", {k: getattr(common, k) for k in ("sigma", "support", "nu", "nplanes")}, "slab", slab, file=sys.stderr)
vis, wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
N = cfg.num_pixels
kw = dict(epsilon=cfg.epsilon, do_wstacking=True, return_info=True)
if getattr(cfg, "fp64", False):
    kw["double_precision_accumulation"] = True
if slab:
    kw["w_slab"] = slab
img = None
for it in range(2):
    del img  # the 32768^2 float64 image is 8 GB and the plane stack takes what is left
    torch.cuda.synchronize(); t = time.perf_counter()
    img, info = wgridder.ms2dirty(d_uvw, d_freq, vis, wgt, N, N, px, px, **kw)
    torch.cuda.synchronize(); dt = time.perf_counter() - t
rep, plan = info["report"], info["plan"]
nvis_run = int(uvw.shape[0] * freq.shape[0])
print(json.dumps({"config": which, "nvis": nvis_run, "nactive": int(info["scan"]["nactive"]), "npix": N, "seconds": dt, "Mvis_s": nvis_run / dt / 1e6,
                  "plan": {k: plan[k] for k in ("sigma", "support", "nu", "nv", "nplanes")},
                  "stage_ms": {k: round(float(v), 2) for k, v in rep.items() if k.startswith("ms_")}, "nblocks": rep.get("nblocks"),
                  "finite": bool(torch.isfinite(img.sum())), "max_mem_GB": torch.cuda.max_memory_allocated() / 1e9}))
