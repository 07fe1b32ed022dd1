#!/usr/bin/env python
"""Development aid: CIP_TRACE timeline of one device-resident and one host-input invert of a slice of the
headline workload.  Usage: CIP_TRACE=1 python tools/trace_invert.py [--times 256]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ska-sdp-continuum-imaging-pipeline_b200"))
import torch  # noqa: E402
from ska_sdp_cip import synthetic as S, wgridder  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--times", type=int, default=256)
a = ap.parse_args()
cfg = S.ska_mid(n_times=a.times)
full = S.ska_mid()
px = S.pixel_size_lm(full, S.make_uvw(full, (0, full.n_times)))
uvw, freq = S.make_uvw(cfg), cfg.frequencies()
dev = torch.device("cuda", 0)
d_vis, d_wgt = S.make_vis_torch(cfg, uvw, freq, px, dev)
d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
N = cfg.num_pixels
for it in range(3):
    print(f"--- device inputs, iteration {it}", file=sys.stderr, flush=True)
    img = wgridder.ms2dirty(d_uvw, d_freq, d_vis, d_wgt, N, N, px, px, epsilon=1e-4, do_wstacking=True)
h_vis, h_wgt = d_vis.cpu().pin_memory().numpy(), d_wgt.cpu().pin_memory().numpy()
h_uvw, h_freq = torch.from_numpy(uvw).pin_memory().numpy(), torch.from_numpy(freq).pin_memory().numpy()
for it in range(3):
    print(f"--- host inputs, iteration {it}", file=sys.stderr, flush=True)
    img = wgridder.ms2dirty(h_uvw, h_freq, h_vis, h_wgt, N, N, px, px, epsilon=1e-4, do_wstacking=True)
