#!/usr/bin/env python
"""Development aid for ncu captures: a few device-resident inverts of a time slice of BASELINE.json configs[1]
(same pixel size, 4096^2 image).  Usage: python tools/run_small_invert.py [--times 64] [--reps 2]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ska-sdp-continuum-imaging-pipeline_b200"))
import torch  # noqa: E402
from ska_sdp_cip import synthetic as S, wgridder  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--times", type=int, default=64)
ap.add_argument("--reps", type=int, default=2)
a = ap.parse_args()
cfg = S.ska_mid(n_times=a.times)
full = S.ska_mid()
px = S.pixel_size_lm(full, S.make_uvw(full, (0, full.n_times)))
uvw, freq = S.make_uvw(cfg), cfg.frequencies()
dev = torch.device("cuda", 0)
d_vis, d_wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
N = cfg.num_pixels
for it in range(a.reps):
    img, info = wgridder.ms2dirty(d_uvw, d_freq, d_vis, d_wgt, N, N, px, px, epsilon=1e-4, do_wstacking=True,
                                  return_info=True)
    print({k: round(v, 3) for k, v in info["report"].items() if k.startswith("ms_")}, flush=True)
