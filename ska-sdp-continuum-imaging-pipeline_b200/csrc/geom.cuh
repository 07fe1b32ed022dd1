// Per-sample geometry shared by the scan, binning and record kernels (device code).
#pragma once

#include "common.h"

namespace cip {

struct SampleGeom {
  int i0, j0, p0;     // first cell (shifted coordinates: zero frequency at nu/2) and first plane
  float fu, fv, fw;   // i0 - left edge of the support, in [0,1)
  double dfu, dfv, dfw;  // the same in float64 (records of the float64 path)
  bool flip;          // w < 0: sample replaced by its Hermitian conjugate
  double w;           // |w| in wavelengths
};

// Sample (row, chan) -> support origin on the oversampled grid and in the w-plane stack.
// Follows the published improved w-stacking scheme (Arras et al. 2021, Ye et al. 2022) that
// ducc0.wgridder.ms2dirty implements (reference call site src/ska_sdp_cip/invert.py:170-183).
// All coordinate arithmetic is float64: grid indices reach 65536 and need ~1e-6 cell accuracy.
__device__ __forceinline__ SampleGeom sample_geom(const GridGeom& g, double um, double vm, double wm, double finv) {
  SampleGeom s;
  double u = um * finv, v = vm * finv, w = wm * finv;
  s.flip = w < 0.0;
  if (s.flip) { u = -u; v = -v; w = -w; }
  if (!g.do_wstacking) w = 0.0;
  s.w = w;
  double xu = u * g.px;
  xu -= rint(xu);
  double xv = v * g.py;
  xv -= rint(xv);
  double us = xu * g.nu + 0.5 * g.nu - 0.5 * g.W;  // left edge of the support, shifted coordinates
  double vs = xv * g.nv + 0.5 * g.nv - 0.5 * g.W;
  double ci = ceil(us), cj = ceil(vs);
  s.i0 = (int)ci;
  s.j0 = (int)cj;
  s.dfu = ci - us;
  s.dfv = cj - vs;
  s.fu = (float)s.dfu;
  s.fv = (float)s.dfv;
  if (g.do_wstacking) {
    double ps = (w - g.w0) * g.inv_dw - 0.5 * g.W;
    double cp = ceil(ps);
    int p0 = (int)cp;
    // samples outside the scanned w range are inactive (zero weight); keep their window legal
    int pmax = g.nplanes - g.W;
    if (p0 < 0 || p0 > pmax) {
      p0 = p0 < 0 ? 0 : pmax;
      cp = (double)p0;
    }
    s.p0 = p0;
    s.dfw = fmin(fmax(cp - ps, 0.0), 1.0);
    s.fw = (float)s.dfw;
  } else {
    s.p0 = 0;
    s.fw = 0.0f;
    s.dfw = 0.0;
  }
  return s;
}

__device__ __forceinline__ uint64_t bin_key(const GridGeom& g, const SampleGeom& s) {
  int tu = (s.i0 + kHalo) / kTile, tv = (s.j0 + kHalo) / kTile;
  int win = g.W <= kFastSupport ? s.p0 / g.win_step : 0;
  return ((uint64_t)win * g.ntu + tu) * g.ntv + tv;
}

// Position of the sample's register window (8 x 8 cells, origin snapped to multiples of 9-W cells,
// see grid.cu) inside its tile: 4 bits per axis.
constexpr int kSubKeyBits = 8;
__device__ __forceinline__ uint32_t window_lattice(const GridGeom& g, const SampleGeom& s) {
  if (g.W > kFastSupport) return 0;
  const int snap = 9 - g.W;
  const int iu = (s.i0 + kHalo) % kTile, iv = (s.j0 + kHalo) % kTile;
  return (uint32_t)((iu / snap) << 4 | (iv / snap));
}

__device__ __forceinline__ bool sample_active(const float2 v, const float* wgt, const uint8_t* mask, int64_t k,
                                              float* wout) {
  float w = wgt ? wgt[k] : 1.0f;
  if (mask && mask[k] == 0) w = 0.0f;
  *wout = w;
  return w != 0.0f && (v.x != 0.0f || v.y != 0.0f);
}

// ES kernel phi(x) = exp(betaW (sqrt(1-x^2) - 1)) for |x| < 1, else 0.
__device__ __forceinline__ float es_kernel(float x, float beta_w) {
  float t = 1.0f - x * x;
  return t >= 0.0f ? expf(beta_w * (sqrtf(t) - 1.0f)) : 0.0f;
}

__device__ __forceinline__ double es_kernel(double x, double beta_w) {
  double t = 1.0 - x * x;
  return t >= 0.0 ? exp(beta_w * (sqrt(t) - 1.0)) : 0.0;
}

__device__ __forceinline__ uint32_t to_fix32(double f) {
  double v = f * kFix32;
  return v >= 4294967295.0 ? 0xffffffffu : (uint32_t)v;
}
__device__ __forceinline__ double from_fix32(uint32_t v) { return (double)v * (1.0 / kFix32); }

}  // namespace cip
