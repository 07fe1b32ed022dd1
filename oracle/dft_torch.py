"""ORACLE (test infrastructure): the float64 direct DFT of oracle/dft.py at sampled pixels, evaluated with
torch so that it can run on the GPU box's device for the BASELINE-size parity checks (1e8 .. 1e9 samples x
~1000 pixels is minutes on host cores, seconds in float64 on the device).  Same definition as dft.dft_dirty
(the quantity ducc0.wgridder.ms2dirty, reference call site src/ska_sdp_cip/invert.py:170-183,
approximates to relative L2 accuracy epsilon); pinned to dft.py by tests/test_oracle.py.
PARITY UNPINNED against ducc0 itself (not installable here).  Only tests/ and bench.py's parity leg import it.
"""
from __future__ import annotations

import numpy as np

SPEED_OF_LIGHT = 299792458.0


def dft_pixels(uvw, freq, vis, wgt, nx, ny, px, py, pixels, *, device=None, rows_per_block=None, row_step=1):
    """D[p] = (1/n_p) sum_k Re(w_k V_k exp(+2 pi i (u_k l_p + v_k m_p - w_k (n_p - 1)))) at the flat pixel indices
    ``pixels`` of an (nx, ny) image, float64.  Inputs may be numpy arrays or torch tensors (any device);
    ``row_step`` > 1 evaluates every row_step-th row only (the caller grids the same subset: the image is
    linear in the samples)."""
    import torch
    if device is None:
        device = vis.device if isinstance(vis, torch.Tensor) else torch.device("cpu")
    as_t = lambda a, dt: (a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))).to(device=device, dtype=dt)
    uvw_t = as_t(uvw, torch.float64)[::row_step]
    finv = as_t(freq, torch.float64) / SPEED_OF_LIGHT
    pix = np.asarray(pixels, dtype=np.int64)
    x = (pix // ny - nx // 2) * float(px)
    y = (pix % ny - ny // 2) * float(py)
    e = x * x + y * y
    nm1 = -e / (np.sqrt(1.0 - e) + 1.0)
    lmn = torch.from_numpy(np.stack([x, y, -nm1], axis=0)).to(device)           # (3, P)
    nrow, nchan, npix = uvw_t.shape[0], finv.shape[0], len(pix)
    if rows_per_block is None:
        rows_per_block = max(1, int(2 ** 27 // max(nchan * npix, 1)))            # ~1 GB of float64 phases per block
    out = torch.zeros(npix, dtype=torch.float64, device=device)
    vis_t = vis if isinstance(vis, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(vis))
    wgt_t = None if wgt is None else (wgt if isinstance(wgt, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(wgt)))
    for r0 in range(0, nrow, rows_per_block):
        r1 = min(nrow, r0 + rows_per_block)
        rows = slice(r0 * row_step, r1 * row_step, row_step)
        v = vis_t[rows].to(device=device, dtype=torch.complex128)
        if wgt_t is not None:
            v = v * wgt_t[rows].to(device=device, dtype=torch.float64)
        geo = uvw_t[r0:r1] @ lmn                                                  # (R, P) metres . direction
        ph = geo[:, None, :] * finv[None, :, None]                                # (R, C, P) cycles
        ph = (ph - torch.round(ph)) * (2.0 * np.pi)
        out += torch.einsum("rc,rcp->p", v.real, torch.cos(ph)) - torch.einsum("rc,rcp->p", v.imag, torch.sin(ph))
    return (out / torch.from_numpy(1.0 + nm1).to(device)).cpu().numpy()


def sample_pixels(nx, ny, n_random=1024, seed=0, always=()):
    """Flat indices of ``always`` (e.g. the injected sources) followed by ``n_random`` random pixels."""
    rng = np.random.default_rng(seed)
    rnd = rng.integers(0, nx * ny, n_random)
    return np.concatenate([np.asarray(always, dtype=np.int64).reshape(-1), rnd]).astype(np.int64)
