"""Deterministic synthetic SKA-shaped visibilities for tests and bench.

The reference ships one real MeasurementSet for its tests (tests/data/mkt_ecdfs25_nano.zip,
74,214 rows x 4 channels x 4 pols), which is not available here; ``meerkat_like`` builds a
stand-in of identical shape.  ``ska_mid`` / ``ska_low`` build the BASELINE.json configurations.
Geometry (uvw, frequencies) is generated with numpy on the host; visibility values are a handful
of point sources plus noise and can be generated with numpy (small cases) or torch (on the GPU,
large cases).  This module is input generation only -- no imaging arithmetic.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional

import numpy as np

SPEED_OF_LIGHT = 299792458.0
SIDEREAL_DAY_S = 86164.0905


@dataclass
class SyntheticConfig:
    name: str
    n_ant: int
    r_min: float
    r_max: float
    latitude_deg: float
    declination_deg: float
    n_times: int
    dt_s: float
    freq0_hz: float
    dfreq_hz: float
    n_chan: int
    num_pixels: int
    seed: int
    h0_deg: float = 0.0
    autos: bool = False
    pixel_margin: float = 0.4   # pixel size = margin / max(|u|,|v|) in wavelengths
    epsilon: float = 1e-4
    fp64: bool = False

    @property
    def n_baselines(self) -> int:
        return self.n_ant * (self.n_ant + 1) // 2 if self.autos else self.n_ant * (self.n_ant - 1) // 2

    @property
    def n_rows(self) -> int:
        return self.n_baselines * self.n_times

    @property
    def n_vis(self) -> int:
        return self.n_rows * self.n_chan

    def frequencies(self) -> np.ndarray:
        return self.freq0_hz + self.dfreq_hz * np.arange(self.n_chan, dtype=np.float64)


def ska_mid(n_times=256, n_chan=64, num_pixels=4096, dfreq_hz=1.0e6, name="ska-mid-aastar") -> SyntheticConfig:
    """BASELINE.json configs[1] (defaults) and configs[3] (n_times=512, n_chan=256, 16384, 0.5 MHz)."""
    return SyntheticConfig(name, 144, 30.0, 75.0e3, -30.7, -45.0, n_times, 8.0, 1.30e9, dfreq_hz, n_chan,
                           num_pixels, 20240)


def ska_low(n_times=32, n_chan=128, num_pixels=8192, high_w=False, name="ska-low") -> SyntheticConfig:
    """configs[2] (defaults) and configs[4] (high_w=True, 32768, epsilon 1e-5, fp64)."""
    if high_w:
        return SyntheticConfig(name + "-high-w", 512, 20.0, 32.5e3, -26.8, 10.0, n_times, 8.0, 150e6, 100e3, n_chan,
                               num_pixels, 20241, h0_deg=45.0, pixel_margin=0.2, epsilon=1e-5, fp64=True)
    return SyntheticConfig(name, 512, 20.0, 32.5e3, -26.8, -26.8, n_times, 8.0, 150e6, 100e3, n_chan, num_pixels, 20241)


def meerkat_like(num_pixels=1024) -> SyntheticConfig:
    """Stand-in of the reference's test MS: 38 times x 1,953 baselines (62 antennas incl. autos) = 74,214
    rows, 4 channels from 959.97 MHz in 835.9375 kHz steps (tests/test_measurement_set.py:26-29)."""
    return SyntheticConfig("meerkat-like", 62, 30.0, 4.0e3, -30.7, -28.0, 38, 8.0, 959969726.5625, 835937.5, 4,
                           num_pixels, 20239, autos=True)


def antenna_enu(cfg: SyntheticConfig) -> np.ndarray:
    rng = np.random.default_rng(cfg.seed)
    r = np.exp(rng.uniform(np.log(cfg.r_min), np.log(cfg.r_max), cfg.n_ant))
    th = rng.uniform(0.0, 2.0 * np.pi, cfg.n_ant)
    return np.stack([r * np.cos(th), r * np.sin(th), np.zeros(cfg.n_ant)], axis=1)


def make_uvw(cfg: SyntheticConfig, time_range: Optional[tuple[int, int]] = None) -> np.ndarray:
    """(n_rows, 3) float64 metres, rows time-major and baseline-minor."""
    enu = antenna_enu(cfg)
    lat = np.radians(cfg.latitude_deg)
    e, n, u = enu[:, 0], enu[:, 1], enu[:, 2]
    xyz = np.stack([-np.sin(lat) * n + np.cos(lat) * u, e, np.cos(lat) * n + np.sin(lat) * u], axis=1)
    i, j = np.triu_indices(cfg.n_ant, k=0 if cfg.autos else 1)
    bl = xyz[j] - xyz[i]
    t0, t1 = time_range or (0, cfg.n_times)
    t = np.arange(t0, t1, dtype=np.float64)
    ha = np.radians(cfg.h0_deg) + (t - cfg.n_times / 2) * cfg.dt_s * 2.0 * np.pi / SIDEREAL_DAY_S
    dec = np.radians(cfg.declination_deg)
    sh, ch = np.sin(ha)[:, None], np.cos(ha)[:, None]
    lx, ly, lz = bl[None, :, 0], bl[None, :, 1], bl[None, :, 2]
    uu = sh * lx + ch * ly
    vv = -np.sin(dec) * ch * lx + np.sin(dec) * sh * ly + np.cos(dec) * lz
    ww = np.cos(dec) * ch * lx - np.cos(dec) * sh * ly + np.sin(dec) * lz
    return np.ascontiguousarray(np.stack([uu, vv, ww], axis=2).reshape(-1, 3))


def pixel_size_lm(cfg: SyntheticConfig, uvw: np.ndarray) -> float:
    """Direction-cosine pixel size that keeps max(|u|,|v|)*pixel = cfg.pixel_margin."""
    fmax = cfg.frequencies().max()
    return float(cfg.pixel_margin / (np.abs(uvw[:, :2]).max() * fmax / SPEED_OF_LIGHT))


def lm_to_asec(pix_lm: float) -> float:
    return float(np.degrees(np.arcsin(pix_lm)) * 3600.0)


def source_list(cfg: SyntheticConfig, pix_lm: float, n_src: int = 8):
    """Fixed point sources at pixel centres: (l, m, flux)."""
    rng = np.random.default_rng(cfg.seed + 17)
    half = cfg.num_pixels // 2
    pi = rng.integers(-int(0.4 * half), int(0.4 * half), n_src)
    pj = rng.integers(-int(0.4 * half), int(0.4 * half), n_src)
    flux = rng.uniform(0.5, 2.0, n_src)
    return pi * pix_lm, pj * pix_lm, flux, pi + half, pj + half


def make_vis_numpy(cfg, uvw, freq, pix_lm, seed=1234, noise=1.0, flag_fraction=0.02):
    """complex64 (nrow,nchan) visibilities, float32 weights with flags folded in (small cases)."""
    l, m, flux, _, _ = source_list(cfg, pix_lm)
    rng = np.random.default_rng(seed)
    nrow, nchan = len(uvw), len(freq)
    finv = freq / SPEED_OF_LIGHT
    vis = np.zeros((nrow, nchan), dtype=np.complex128)
    for ls, ms_, fs in zip(l, m, flux):
        nm1 = -(ls * ls + ms_ * ms_) / (np.sqrt(1.0 - ls * ls - ms_ * ms_) + 1.0)
        ph = (uvw[:, 0] * ls + uvw[:, 1] * ms_ - uvw[:, 2] * nm1)[:, None] * finv[None, :]  # ms2dirty's w sign
        vis += fs * np.exp(-2j * np.pi * ph)
    vis += noise * (rng.normal(size=vis.shape) + 1j * rng.normal(size=vis.shape))
    wgt = rng.uniform(0.5, 1.5, size=vis.shape).astype(np.float32)
    wgt[rng.random(vis.shape) < flag_fraction] = 0.0
    return vis.astype(np.complex64), wgt


def make_vis_torch(cfg, uvw, freq, pix_lm, device, seed=1234, noise=1.0, flag_fraction=0.02, rows_per_block=1 << 16):
    """Same model generated on the GPU with torch (large cases); returns CUDA tensors (vis c64, wgt f32)."""
    import torch
    l, m, flux, _, _ = source_list(cfg, pix_lm)
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    nrow, nchan = len(uvw), len(freq)
    d_uvw = torch.from_numpy(uvw).to(device)
    finv = torch.from_numpy(freq / SPEED_OF_LIGHT).to(device)
    vis = torch.empty((nrow, nchan), dtype=torch.complex64, device=device)
    wgt = torch.empty((nrow, nchan), dtype=torch.float32, device=device)
    for r0 in range(0, nrow, rows_per_block):
        r1 = min(nrow, r0 + rows_per_block)
        blk = torch.zeros((r1 - r0, nchan), dtype=torch.complex128, device=device)
        for ls, ms_, fs in zip(l, m, flux):
            nm1 = -(ls * ls + ms_ * ms_) / (np.sqrt(1.0 - ls * ls - ms_ * ms_) + 1.0)
            ph = (d_uvw[r0:r1, 0] * ls + d_uvw[r0:r1, 1] * ms_ - d_uvw[r0:r1, 2] * nm1)[:, None] * finv[None, :]
            ph = (ph - torch.round(ph)) * (-2.0 * np.pi)
            blk += fs * torch.complex(torch.cos(ph), torch.sin(ph))
        nz = torch.randn((r1 - r0, nchan, 2), generator=gen, device=device, dtype=torch.float32) * noise
        vis[r0:r1] = blk.to(torch.complex64) + torch.view_as_complex(nz)
        w = torch.rand((r1 - r0, nchan), generator=gen, device=device, dtype=torch.float32) + 0.5
        fl = torch.rand((r1 - r0, nchan), generator=gen, device=device, dtype=torch.float32) < flag_fraction
        wgt[r0:r1] = torch.where(fl, torch.zeros_like(w), w)
    return vis, wgt


def four_pol(vis_i: np.ndarray, wgt_i: np.ndarray, seed=99):
    """Expand Stokes-I data to the (nrow, nchan, 4) layout a MeasurementSet reader returns:
    XX = I + q, YY = I - q (so (XX+YY)/2 = I), cross hands noise; per-hand weights whose Stokes-I
    combination 4/(1/wxx+1/wyy) is wgt_i where unflagged; zero effective weight becomes a flag."""
    rng = np.random.default_rng(seed)
    q = (0.1 * (rng.normal(size=vis_i.shape) + 1j * rng.normal(size=vis_i.shape))).astype(np.complex64)
    vis = np.empty(vis_i.shape + (4,), dtype=np.complex64)
    vis[..., 0] = vis_i + q
    vis[..., 3] = vis_i - q
    vis[..., 1] = q
    vis[..., 2] = np.conj(q)
    flags = np.zeros(vis.shape, dtype=bool)
    flagged = wgt_i == 0
    flags[..., 0] = flagged & (rng.random(vis_i.shape) < 0.5)
    flags[..., 3] = flagged & ~flags[..., 0]
    w = np.where(flagged, 1.0, wgt_i).astype(np.float32)
    wgt = np.repeat((0.5 * w)[..., None], 4, axis=-1).astype(np.float32)  # 4/(2/w'+2/w') with w' = w/2 -> w
    return vis, flags, wgt
