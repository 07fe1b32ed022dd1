#!/usr/bin/env python
"""Offline design of the exponential-of-semicircle (ES) gridding kernel table.

For support W (grid cells) and oversampling sigma the 1-D kernel is
    phi(x) = exp(beta*W*(sqrt(1-x^2)-1)),  |x|<1   (x = 2t/W)
Gridding followed by division with the kernel transform c(xi) leaves the
aliasing error (Poisson summation)
    err(xi)^2 = sum_{k!=0} c(xi+k)^2 / c(xi)^2 ,   |xi| <= 1/(2 sigma)
This script picks, per (W, sigma), the beta that minimises the RMS of err over
the retained image band and prints the table that csrc/cip_plan.cpp and
oracle/wstack.py carry.  It is a design-time tool, not part of the product path.
"""
import numpy as np
from numpy.polynomial.legendre import leggauss
from scipy.optimize import minimize_scalar

_GL = leggauss(3000)

def chat(xi, W, beta):
    # x = sin(theta) removes the square-root end-point singularity of phi
    th, w = _GL
    th = th * (np.pi / 2)
    w = w * (np.pi / 2)
    x = np.sin(th)
    phi = np.exp(beta * W * (np.cos(th) - 1.0)) * np.cos(th)
    t = x * (W / 2.0)
    return (W / 2.0) * np.sum(w * phi * np.cos(2 * np.pi * np.outer(xi, t)), axis=1)

def rms_err(W, sigma, beta, nxi=65, kmax=8):
    xi = np.linspace(0.0, 0.5 / sigma, nxi)
    c0 = chat(xi, W, beta)
    s = np.zeros_like(xi)
    for k in range(1, kmax + 1):
        s += chat(xi + k, W, beta) ** 2 + chat(xi - k, W, beta) ** 2
    e = np.sqrt(s) / np.abs(c0)
    return np.sqrt(np.mean(e ** 2)), e.max()

if __name__ == "__main__":
    print("# W sigma beta rms_err max_err")
    for sigma in (1.25, 1.5, 1.75, 2.0):
        for W in range(4, 17):
            f = lambda b: np.log(rms_err(W, sigma, b)[0])
            b0 = np.pi * (1 - 1 / (2 * sigma))
            r = minimize_scalar(f, bounds=(0.7 * b0, 1.15 * b0), method="bounded", options={"xatol": 1e-4})
            rms, mx = rms_err(W, sigma, r.x)
            print(f"{W:2d} {sigma:.2f} {r.x:.5f} {rms:.3e} {mx:.3e}   (beta/b0={r.x/b0:.4f})")
