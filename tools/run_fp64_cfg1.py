"""configs[1] input through the float64 path (complex128 grids, epsilon 1e-5): stage times of the correctness-path
gridders, for DESIGN.md.  python tools/run_fp64_cfg1.py [n_times]"""
import json, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "ska-sdp-continuum-imaging-pipeline_b200"))
import numpy as np, torch
from ska_sdp_cip import synthetic as S, wgridder

n_times = int(sys.argv[1]) if len(sys.argv) > 1 else 256
cfg = S.ska_mid(n_times=n_times)
full = S.ska_mid(n_times=256)
dev = torch.device("cuda", 0)
px = S.pixel_size_lm(full, S.make_uvw(full))
uvw, freq = S.make_uvw(full, (0, n_times)), cfg.frequencies()
vis, wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
N = cfg.num_pixels
for eps, fp64 in ((1e-4, True), (1e-5, True), (1e-5, False)):
    for it in range(2):
        torch.cuda.synchronize(); t = time.perf_counter()
        img, info = wgridder.ms2dirty(d_uvw, d_freq, vis, wgt, N, N, px, px, epsilon=eps, do_wstacking=True, return_info=True,
                                      double_precision_accumulation=fp64)
        torch.cuda.synchronize(); dt = time.perf_counter() - t
    rep, plan = info["report"], info["plan"]
    print(json.dumps({"epsilon": eps, "fp64": fp64, "nvis": int(uvw.shape[0] * freq.shape[0]), "seconds": round(dt, 4),
                      "plan": {k: plan[k] for k in ("sigma", "support", "nu", "nplanes")},
                      "stage_ms": {k: round(float(v), 2) for k, v in rep.items() if k.startswith("ms_")}}), flush=True)
    del img
