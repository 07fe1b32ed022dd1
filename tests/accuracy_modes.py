"""Measured accuracy of the float32 gridder variants against the float64 DFT (run on a GPU box):
CIP_GRID_MMA=0 (FFMA2 rank-1 updates), 1 (3xTF32 mma.sync tiles added with float32 additions), 2 (two-term bfloat16
split, m16n8k16) and
CIP_GRID_TC=1 (tcgen05: split operands accumulated in tensor memory, whose float32 adder truncates).  Two problems: the tests' random one (few visibilities per
register window) and a dense one (thousands of visibilities per window, the case that stresses accumulation).
For each, the planner's epsilon = 1e-4 plan and a forced sigma = 2, W = 8 plan (kernel error ~1e-7, so what is
left is the arithmetic)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ska-sdp-continuum-imaging-pipeline_b200"))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import random_problem  # noqa: E402
from oracle import dft  # noqa: E402  (checker only)
from ska_sdp_cip import wgridder  # noqa: E402

problems = {
    "random": random_problem(),
    "dense": random_problem(nrow=4000, nchan=16, nx=128, seed=3, uvscale=3.0, wscale=1.0)[:5] + (2e-4,),
}
for name, (uvw, freq, vis, wgt, nx, px) in problems.items():
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px)
    for plan in ("eps=1e-4", "sigma=2,W=8"):
        if plan != "eps=1e-4":
            os.environ["CIP_SIGMA"], os.environ["CIP_SUPPORT"] = "2.0", "8"
        out = []
        for env in ({"CIP_GRID_MMA": "0"}, {"CIP_GRID_MMA": "1"}, {"CIP_GRID_MMA": os.environ.get("ACC_BF16_MODE", "2")}, {"CIP_GRID_TC": "1"}):
            os.environ.update(env)
            img = wgridder.ms2dirty(uvw, freq, vis, wgt, nx, nx, px, px, epsilon=1e-4, do_wstacking=True)
            for k in env:
                os.environ.pop(k)
            out.append(dft.rel_rms(np.asarray(img), ref))
        os.environ.pop("CIP_SIGMA", None)
        os.environ.pop("CIP_SUPPORT", None)
        print(f"{name:7s} {plan:12s} rel rms error vs DFT: FFMA2 {out[0]:.3e}  3xTF32 mma.sync + float32 adds {out[1]:.3e}  2-term bf16 mma.sync {out[2]:.3e}  tcgen05 / tensor memory {out[3]:.3e}")
