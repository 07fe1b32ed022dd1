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

// Constants of one binning launch derived from the geometry (loop-invariant, a few registers).
struct BinConsts {
  double nu_d, nv_d, off_u, off_v;  // left edge of the support: x * nu + (nu - W) / 2 (shifted coordinates)
  double w_scale, w_off;            // first-plane coordinate: |w| * inv_dw - (w0 * inv_dw + W / 2)
  uint32_t ws_magic;                // ceil(2^32 / win_step): p0 / win_step = umulhi(p0, ws_magic), exact for p0 < 2^29
  uint32_t snap_rcp;                // ceil(256 / snap): i / snap = (i * snap_rcp) >> 8, exact for i < 8, snap <= 5
  int snap, pmax;
  bool fast;                        // W <= kFastSupport: windows and register-window lattice in use
};
// Computed once on the host and passed to the kernels by value: evaluated per thread it cost more instructions (a 64-bit
// division among them) than the two loop iterations a 64-channel row needs.
__host__ __device__ __forceinline__ BinConsts bin_consts(const GridGeom& g) {
  BinConsts k;
  k.nu_d = (double)g.nu;
  k.nv_d = (double)g.nv;
  k.off_u = 0.5 * g.nu - 0.5 * g.W;
  k.off_v = 0.5 * g.nv - 0.5 * g.W;
  k.w_scale = g.inv_dw;
  k.w_off = -g.w0 * g.inv_dw - 0.5 * g.W;
  k.fast = g.W <= kFastSupport;
  k.snap = k.fast ? 9 - g.W : 1;
  k.snap_rcp = (256u + k.snap - 1) / k.snap;
  k.ws_magic = g.win_step > 1 ? (uint32_t)((0x100000000ull + g.win_step - 1) / g.win_step) : 0u;
  k.pmax = g.nplanes - g.W;
  return k;
}

// Per-row part: the Hermitian flip depends only on the sign of w in metres (frequencies are positive),
// and the pixel scale is folded into the row's coordinates.
struct RowGeom {
  double upx, vpy, wabs;
  bool flip;
};
__device__ __forceinline__ RowGeom row_geom(const GridGeom& g, double um, double vm, double wm) {
  RowGeom r;
  r.flip = wm < 0.0;
  const double sg = r.flip ? -1.0 : 1.0;
  r.upx = sg * um * g.px;
  r.vpy = sg * vm * g.py;
  r.wabs = g.do_wstacking ? sg * wm : 0.0;
  return r;
}

// Sample (row, chan) -> support origin on the oversampled grid and in the w-plane stack.
// Follows the published improved w-stacking scheme (Arras et al. 2021, Ye et al. 2022) that
// ducc0.wgridder.ms2dirty implements (reference call site src/ska_sdp_cip/invert.py:170-183).
// All coordinate arithmetic is float64: grid indices reach 65536 and need ~1e-6 cell accuracy.
// finv = frequency / c (tabulated once per call).  FRAC = false skips the fractional offsets (key passes).
// Rounding without the conversion unit: FRND.F64 / F2I.F64 issue at a quarter of the FP64 add rate and the binning passes
// are bound by instruction issue.  x + 1.5 * 2^52 rounds x to the nearest integer (ties to even, exact for |x| < 2^51)
// and leaves it, as a two's-complement int, in the low word of the sum.
constexpr double kRoundMagic = 6755399441055744.0;
__device__ __forceinline__ double rint_magic(double x) { return (x + kRoundMagic) - kRoundMagic; }
// ceil(x) as int (|x| < 2^31) and, in *frac, ceil(x) - x (one rounding, as the direct difference has)
__device__ __forceinline__ int ceil_magic(double x, double* frac) {
  const double t = x + kRoundMagic, r = t - kRoundMagic;  // r = rint(x)
  const bool up = r < x;
  *frac = (r - x) + (up ? 1.0 : 0.0);  // r - x is exact
  return __double2loint(t) + (up ? 1 : 0);
}

// exp(2 pi i ph) in float32 for a phase in turns held in float64 (any magnitude below 2^49): quadrant and remainder by
// magic-number rounding, Taylor polynomials on |x| <= 1/8 turn (errors below 3e-8) -- a third of the instructions of
// rint + sincospif and one conversion instead of four.
__device__ __forceinline__ void sincos_turns(double ph, float* sn, float* cs) {
  const double t = fma(ph, 4.0, kRoundMagic);
  const int q = __double2loint(t);                              // rint(4 ph)
  const float x = (float)fma(t - kRoundMagic, -0.25, ph);       // ph - q / 4, exact, in [-1/8, 1/8]
  const float x2 = x * x;
  float s = fmaf(x2, 42.058693944897655f, -76.70585975306136f);  // (2 pi)^9 / 9!, -(2 pi)^7 / 7!
  s = fmaf(x2, s, 81.60524927607504f);                           // (2 pi)^5 / 5!
  s = fmaf(x2, s, -41.34170224039975f);                          // -(2 pi)^3 / 3!
  s = fmaf(x2, s, 6.283185307179586f);
  s *= x;
  float c = fmaf(x2, -26.42625678337438f, 60.24464137187666f);   // -(2 pi)^10 / 10!, (2 pi)^8 / 8!
  c = fmaf(x2, c, -85.45681720669373f);                          // -(2 pi)^6 / 6!
  c = fmaf(x2, c, 64.93939402266829f);                           // (2 pi)^4 / 4!
  c = fmaf(x2, c, -19.739208802178716f);                         // -(2 pi)^2 / 2
  c = fmaf(x2, c, 1.0f);
  const float a = (q & 1) ? c : s, b = (q & 1) ? s : c;          // quadrant 1, 3: sin <- cos, cos <- -sin
  *sn = (q & 2) ? -a : a;
  *cs = ((q + 1) & 2) ? -b : b;
}

template <bool FRAC>
__device__ __forceinline__ SampleGeom sample_geom(const GridGeom& g, const BinConsts& k, const RowGeom& r, double finv) {
  SampleGeom s;
  s.flip = r.flip;
  double xu = r.upx * finv;
  xu -= rint_magic(xu);
  double xv = r.vpy * finv;
  xv -= rint_magic(xv);
  const double us = fma(xu, k.nu_d, k.off_u), vs = fma(xv, k.nv_d, k.off_v);
  double du, dv;
  s.i0 = ceil_magic(us, &du);
  s.j0 = ceil_magic(vs, &dv);
  s.w = r.wabs * finv;
  if (FRAC) {
    s.dfu = du;
    s.dfv = dv;
    s.fu = (float)s.dfu;
    s.fv = (float)s.dfv;
  }
  if (g.do_wstacking) {
    const double ps = fma(s.w, k.w_scale, k.w_off);
    double dp;
    int p0 = ceil_magic(ps, &dp);
    // samples outside the scanned w range are inactive (zero weight); keep their window legal
    // (their records carry a zero visibility, so the offset they get does not matter)
    if (p0 < 0 || p0 > k.pmax) {
      p0 = p0 < 0 ? 0 : k.pmax;
      dp = 0.0;
    }
    s.p0 = p0;
    if (FRAC) {
      s.dfw = dp;  // in [0, 1) by construction
      s.fw = (float)s.dfw;
    }
  } else {
    s.p0 = 0;
    s.fw = 0.0f;
    s.dfw = 0.0;
  }
  return s;
}

// Decomposition of the support origin into (window, tile, position inside the tile, register-window lattice).
struct BinCell {
  int tu, tv, win;
  uint32_t i0r, j0r, p0r;  // origin inside the tile / first plane inside the window
  uint32_t lu, lv;         // lattice position of the register window (i0r / snap, j0r / snap)
};
__device__ __forceinline__ BinCell bin_cell(const GridGeom& g, const BinConsts& k, const SampleGeom& s) {
  BinCell b;
  const uint32_t iu = (uint32_t)(s.i0 + kHalo), iv = (uint32_t)(s.j0 + kHalo);  // non-negative by construction
  b.tu = (int)(iu / kTile);
  b.tv = (int)(iv / kTile);
  b.i0r = iu % kTile;
  b.j0r = iv % kTile;
  b.win = k.fast && k.ws_magic ? (int)__umulhi((uint32_t)s.p0, k.ws_magic) : (k.fast ? s.p0 : 0);
  b.p0r = (uint32_t)(s.p0 - b.win * g.win_step);
  b.lu = k.fast ? (b.i0r * k.snap_rcp) >> 8 : 0u;
  b.lv = k.fast ? (b.j0r * k.snap_rcp) >> 8 : 0u;
  return b;
}

constexpr uint32_t kNoKey = 0xffffffffu;
// Sort key of a sample: its (window, tile) bin and the register-window lattice position inside the tile, dense;
// kNoKey for samples whose window the current binning pass does not cover.
__device__ __forceinline__ uint32_t bin_key(const GridGeom& g, const BinCell& b) {
  if (b.win < g.win_begin || b.win >= g.win_end) return kNoKey;
  const uint32_t bin = ((uint32_t)(b.win - g.win_begin) * (uint32_t)g.ntu + (uint32_t)b.tu) * (uint32_t)g.ntv + (uint32_t)b.tv;
  return (bin * (uint32_t)g.nlat + b.lu) * (uint32_t)g.nlat + b.lv;
}
// Inverse, for the consumers of a job.
struct KeyParts {
  int win, tu, tv, lu, lv;
};
__device__ __forceinline__ KeyParts key_parts(const GridGeom& g, uint32_t key) {
  KeyParts k;
  k.lv = (int)(key % (uint32_t)g.nlat);
  key /= (uint32_t)g.nlat;
  k.lu = (int)(key % (uint32_t)g.nlat);
  key /= (uint32_t)g.nlat;
  k.tv = (int)(key % (uint32_t)g.ntv);
  key /= (uint32_t)g.ntv;
  k.tu = (int)(key % (uint32_t)g.ntu);
  k.win = (int)(key / (uint32_t)g.ntu) + g.win_begin;
  return k;
}

// Weight (mask applied) and visibility of the sample at element k of wgt / mask resp. vis, in the input precision.
__device__ __forceinline__ double sample_weight(const DevInputs& in, int64_t k) {
  double w = 1.0;
  if (in.wgt) w = in.wide ? reinterpret_cast<const double*>(in.wgt)[k] : (double)in.wgt[k];
  if (in.mask && in.mask[k] == 0) w = 0.0;
  return w;
}

// Channels of a row and where its samples live: sample (row, c) is element vbase + c of vis and wbase + c of
// wgt / mask (dense rows: all channels at row * stride; jagged rows: see DevInputs).
struct RowSpan {
  int c_begin, c_end;
  int64_t vbase, wbase;
};
__device__ __forceinline__ RowSpan row_span(const DevInputs& in, int64_t row) {
  RowSpan s;
  if (in.voff) {
    s.c_begin = in.ch0[row];
    s.c_end = in.ch1[row];
    s.vbase = s.wbase = in.voff[row] - s.c_begin;
  } else {
    s.c_begin = 0;
    s.c_end = (int)in.nchan;
    s.vbase = row * in.vstride;
    s.wbase = row * in.wstride;
  }
  return s;
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

__device__ __forceinline__ uint64_t to_fix18(double f) {  // f in [0, 1]: round(f 2^18), without the conversion unit
  const uint32_t q = (uint32_t)__double2loint(fma(f, kFix18, kRoundMagic));
  return q > 262143u ? 262143u : q;
}
__device__ __forceinline__ uint32_t to_fix32(double f) {
  double v = f * kFix32;
  return v >= 4294967295.0 ? 0xffffffffu : (uint32_t)v;
}
__device__ __forceinline__ double from_fix32(uint32_t v) { return (double)v * (1.0 / kFix32); }

}  // namespace cip
