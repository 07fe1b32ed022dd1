"""ORACLE (test infrastructure): restatement of the gridder planning step (G1).

ducc0 chooses (support W, oversampling sigma) from epsilon with a cost model and
derives grid size, w-plane spacing and plane count from the image geometry
(call site src/ska_sdp_cip/invert.py:170-183).  ducc0's
tuned kernel table is not available here, so the product carries its own table
of plain exponential-of-semicircle kernels designed by tools/kernel_params.py;
this module restates the product's planning rules independently in Python so
tests can check ``cip_plan_query`` against it.  PARITY UNPINNED against ducc0.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

# (W -> (beta, rms aliasing error)) per oversampling factor; output of tools/kernel_params.py
ES_TABLE = {
    1.25: {4: (1.68962, 5.411e-03), 5: (1.75919, 1.388e-03), 6: (1.79517, 3.589e-04), 7: (1.81710, 9.384e-05),
           8: (1.83147, 2.436e-05), 9: (1.84139, 6.284e-06), 10: (1.84861, 1.611e-06), 11: (1.85416, 4.097e-07),
           12: (1.85851, 1.029e-07), 13: (1.81293, 3.200e-08), 14: (1.82197, 7.831e-09), 15: (1.82927, 1.908e-09),
           16: (1.83537, 4.649e-10)},
    1.5: {4: (1.92876, 1.721e-03), 5: (1.98510, 2.928e-04), 6: (2.01566, 5.013e-05), 7: (2.03459, 8.651e-06),
          8: (2.04697, 1.485e-06), 9: (2.05564, 2.542e-07), 10: (2.06202, 4.336e-08), 11: (2.00871, 8.418e-09),
          12: (2.02120, 1.373e-09), 13: (2.03091, 2.240e-10), 14: (2.03867, 3.673e-11), 15: (2.04508, 6.021e-12),
          16: (2.05032, 9.879e-13)},
    1.75: {4: (2.09316, 8.551e-04), 5: (2.14372, 1.141e-04), 6: (2.17179, 1.540e-05), 7: (2.18902, 2.079e-06),
           8: (2.20031, 2.800e-07), 9: (2.20827, 3.753e-08), 10: (2.15092, 5.730e-09), 11: (2.16563, 7.293e-10),
           12: (2.17685, 9.343e-11), 13: (2.18569, 1.201e-11), 14: (2.19284, 1.547e-12), 15: (2.19860, 1.997e-13),
           16: (2.20328, 2.666e-14)},
    2.0: {4: (2.21529, 5.397e-04), 5: (2.26213, 6.111e-05), 6: (2.28840, 7.002e-06), 7: (2.30448, 7.985e-07),
          8: (2.31509, 9.086e-08), 9: (2.32256, 1.031e-08), 10: (2.26900, 1.300e-09), 11: (2.28272, 1.408e-10),
          12: (2.33548, 1.471e-11), 13: (2.30147, 1.672e-12), 14: (2.30810, 1.827e-13), 15: (2.31335, 2.023e-14),
          16: (2.34320, 3.786e-15)},
}

# cost-model constants (seconds per unit); see DESIGN.md "planning"
COST_VIS_FAST = 6.0e-11      # one visibility through the residue-owner gridder (W <= 8)
COST_CELL_UPDATE = 2.0e-12   # one complex cell update in the wide-support gridder (W > 8)
COST_GRID_CELL = 1.0e-11     # zero + FFT + w-screen of one complex grid cell of one plane
FAST_SUPPORT = 8


@dataclass
class Plan:
    nx: int
    ny: int
    px: float
    py: float
    sigma: float
    W: int
    beta: float
    nu: int
    nv: int
    nshift: float
    dw: float
    w0: float
    nplanes: int
    nm1_min: float


def good_size(n: int) -> int:
    """Smallest even 7-smooth integer >= n."""
    n = max(int(n), 2)
    while True:
        m = n
        if m % 2 == 0:
            for p in (2, 3, 5, 7):
                while m % p == 0:
                    m //= p
            if m == 1:
                return n
        n += 1


def nm1_min_of(nx, ny, px, py) -> float:
    l = (nx // 2) * px
    m = (ny // 2) * py
    e = l * l + m * m
    if e >= 1.0:
        return -1.0
    return -e / (math.sqrt(1.0 - e) + 1.0)


def make_plan(nx, ny, px, py, epsilon, wmin_d, wmax_d, nvis, *, fp64=False, do_wstacking=True,
              sigma=None, W=None) -> Plan:
    eps1 = epsilon / (30.0 if fp64 else 3.0) if do_wstacking else epsilon / (20.0 if fp64 else 2.0)
    nm1_min = nm1_min_of(nx, ny, px, py) if do_wstacking else 0.0
    nshift = -0.5 * nm1_min
    half = 0.5 * abs(nm1_min)
    scale = 2.0 if fp64 else 1.0
    best = None
    for sg in (sigma,) if sigma else (1.25, 1.5, 1.75, 2.0):
        if not sigma and not fp64 and sg < 1.4:
            continue  # fp32 grids: keep the 1/c(xi) edge gain modest
        tab = ES_TABLE[sg]
        if W:
            w = W
        else:
            ok = [k for k in sorted(tab) if tab[k][1] <= eps1]
            if not ok:
                continue
            w = ok[0]
        nu = good_size(math.ceil(nx * sg))
        nv = good_size(math.ceil(ny * sg))
        if not do_wstacking or half == 0.0:
            dw, npl = 1.0, 1
        else:
            dw = 1.0 / (2.0 * sg * half)
            npl = int(math.floor((wmax_d - wmin_d) / dw + w))
        wfac = w if do_wstacking else 1
        per_vis = COST_VIS_FAST if w <= FAST_SUPPORT else w * w * wfac * COST_CELL_UPDATE
        cost = scale * (nvis * per_vis + npl * nu * nv * COST_GRID_CELL)
        if best is None or cost < best[0]:
            best = (cost, sg, w, nu, nv, dw, npl)
    if best is None:
        raise ValueError("epsilon too small for the kernel table")
    _, sg, w, nu, nv, dw, npl = best
    if do_wstacking and half != 0.0:
        w0 = 0.5 * (wmin_d + wmax_d) - 0.5 * (npl - 1) * dw
    else:
        w0 = 0.0
    return Plan(nx, ny, px, py, sg, w, ES_TABLE[sg][w][0], nu, nv, nshift, dw, w0, npl, nm1_min)
