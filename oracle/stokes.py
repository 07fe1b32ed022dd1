"""ORACLE (test infrastructure): numpy restatement of the reference's pre/post arithmetic.

P1  StokesIGridderInput.from_measurement_set_reader   src/ska_sdp_cip/invert.py:78-116
P2  StokesIGridderInput.effective_weights             src/ska_sdp_cip/invert.py:72-76
P3  pixel size conversion, weight sum                 src/ska_sdp_cip/invert.py:163,184
P4  final 1/sum(weights) scaling                      src/ska_sdp_cip/invert.py:149
P5  integrate_weighted_images                         src/ska_sdp_cip/invert.py:200-209
PINNED: tests/golden/stokes_small.npz was produced by the reference's own invert.py imported with
stubbed third-party modules (tools/make_golden.py); tests/test_oracle.py compares bit for bit.
"""
from __future__ import annotations

import warnings

import numpy as np


def stokes_i(vis4, flags4, weights4):
    """(vis_I complex64, flags_I bool, weights_I float32), dtypes as numpy produces them in the reference."""
    vis_i = 0.5 * (vis4[..., 0] + vis4[..., 3])
    flags_i = flags4[..., (0, 3)].max(axis=-1)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with np.errstate(divide="ignore"):
            w_i = 4.0 / (1.0 / weights4[..., 0] + 1.0 / weights4[..., 3])
    return vis_i, flags_i, w_i


def effective_weights(flags_i, weights_i):
    return np.logical_not(flags_i) * weights_i


def pixel_size_lm(pixel_size_asec: float) -> float:
    return float(np.sin(np.radians(pixel_size_asec / 3600.0)))


def integrate_weighted_images(weighted_images):
    images, weights = zip(*weighted_images)
    return sum(images) / sum(weights)
