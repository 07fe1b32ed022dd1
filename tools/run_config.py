"""Run one BASELINE.json configuration (per-GPU share) through ms2dirty on device inputs and print the plan,
the stage times and the memory high-water mark:  python tools/run_config.py cfg4|cfg3|cfg5 [n_times]"""
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
else:                  # configs[4] share: high-w SKA-Low, fp64, eps 1e-5
    cfg = S.ska_low(n_times=int(sys.argv[2]) if len(sys.argv) > 2 else 4, num_pixels=int(sys.argv[3]) if len(sys.argv) > 3 else 8192, high_w=True)
    full = cfg
dev = torch.device("cuda", 0)
uvw_full = S.make_uvw(full, (0, full.n_times))
px = S.pixel_size_lm(full, uvw_full)
uvw = S.make_uvw(full, (0, cfg.n_times))
freq = cfg.frequencies()
vis, wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
N = cfg.num_pixels
kw = dict(epsilon=cfg.epsilon, do_wstacking=True, return_info=True)
if getattr(cfg, "fp64", False):
    kw["double_precision_accumulation"] = True
for it in range(2):
    torch.cuda.synchronize(); t = time.perf_counter()
    img, info = wgridder.ms2dirty(d_uvw, d_freq, vis, wgt, N, N, px, px, **kw)
    torch.cuda.synchronize(); dt = time.perf_counter() - t
rep, plan = info["report"], info["plan"]
print(json.dumps({"config": which, "nvis": int(cfg.n_vis), "npix": N, "seconds": dt, "Mvis_s": cfg.n_vis / dt / 1e6,
                  "plan": {k: plan[k] for k in ("sigma", "support", "nu", "nv", "nplanes")},
                  "stage_ms": {k: round(float(v), 2) for k, v in rep.items() if k.startswith("ms_")}, "nblocks": rep.get("nblocks"),
                  "finite": bool(torch.isfinite(img).all()), "max_mem_GB": torch.cuda.max_memory_allocated() / 1e9}))
