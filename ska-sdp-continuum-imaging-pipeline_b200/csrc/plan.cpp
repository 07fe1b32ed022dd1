// Host-side planning of the w-stacking gridder (stage G1 of SURVEY.md section 8a): choice of
// oversampling and kernel support from epsilon, grid size, w-plane layout, and the
// Fourier transform of the gridding kernel used by the final correction.
//
// Replaces the planning that happens inside ducc0.wgridder.ms2dirty (reference call site
// src/ska_sdp_cip/invert.py:170-183).  ducc0's tuned kernel database is not available, so the
// library carries its own table of exponential-of-semicircle kernels
//     phi(x) = exp(beta*W*(sqrt(1-x^2)-1)),  |x| < 1,
// designed offline by tools/kernel_params.py (beta minimises the rms aliasing error over the
// retained image band for each (W, sigma)).
#include <algorithm>
#include <cmath>
#include <cstring>

#include "common.h"

namespace cip {

namespace {

struct TableRow { double sigma; KernelEntry e[kMaxSupport - kMinSupport + 1]; };

// (beta, rms aliasing error) for W = 4..16; output of tools/kernel_params.py
const TableRow kTable[] = {
    {1.25, {{1.68962, 5.411e-03}, {1.75919, 1.388e-03}, {1.79517, 3.589e-04}, {1.81710, 9.384e-05},
            {1.83147, 2.436e-05}, {1.84139, 6.284e-06}, {1.84861, 1.611e-06}, {1.85416, 4.097e-07},
            {1.85851, 1.029e-07}, {1.81293, 3.200e-08}, {1.82197, 7.831e-09}, {1.82927, 1.908e-09},
            {1.83537, 4.649e-10}}},
    {1.5, {{1.92876, 1.721e-03}, {1.98510, 2.928e-04}, {2.01566, 5.013e-05}, {2.03459, 8.651e-06},
           {2.04697, 1.485e-06}, {2.05564, 2.542e-07}, {2.06202, 4.336e-08}, {2.00871, 8.418e-09},
           {2.02120, 1.373e-09}, {2.03091, 2.240e-10}, {2.03867, 3.673e-11}, {2.04508, 6.021e-12},
           {2.05032, 9.879e-13}}},
    {1.75, {{2.09316, 8.551e-04}, {2.14372, 1.141e-04}, {2.17179, 1.540e-05}, {2.18902, 2.079e-06},
            {2.20031, 2.800e-07}, {2.20827, 3.753e-08}, {2.15092, 5.730e-09}, {2.16563, 7.293e-10},
            {2.17685, 9.343e-11}, {2.18569, 1.201e-11}, {2.19284, 1.547e-12}, {2.19860, 1.997e-13},
            {2.20328, 2.666e-14}}},
    {2.0, {{2.21529, 5.397e-04}, {2.26213, 6.111e-05}, {2.28840, 7.002e-06}, {2.30448, 7.985e-07},
           {2.31509, 9.086e-08}, {2.32256, 1.031e-08}, {2.26900, 1.300e-09}, {2.28272, 1.408e-10},
           {2.33548, 1.471e-11}, {2.30147, 1.672e-12}, {2.30810, 1.827e-13}, {2.31335, 2.023e-14},
           {2.34320, 3.786e-15}}},
};
constexpr int kNumSigma = sizeof(kTable) / sizeof(kTable[0]);

// Cost model (seconds per unit on one B200; calibrated from bench runs, see DESIGN.md "planning").
constexpr double kCostVisFast = 6.0e-11;    // one visibility through the residue-owner gridder (W <= 8)
constexpr double kCostCellUpdate = 2.0e-12; // one complex cell update in the wide-support gridder (W > 8)
constexpr double kCostGridCell = 1.0e-11;   // zero + FFT + w-screen of one complex cell of one plane

}  // namespace

int good_size(int n) {
  n = std::max(n, 2);
  for (;; ++n) {
    if (n % 2) continue;
    int m = n;
    for (int p : {2, 3, 5, 7})
      while (m % p == 0) m /= p;
    if (m == 1) return n;
  }
}

bool kernel_lookup(double sigma, int W, KernelEntry* out) {
  if (W < kMinSupport || W > kMaxSupport) return false;
  for (int s = 0; s < kNumSigma; ++s)
    if (std::fabs(kTable[s].sigma - sigma) < 1e-9) {
      *out = kTable[s].e[W - kMinSupport];
      return true;
    }
  return false;
}

// Gauss-Legendre nodes and weights on [-1, 1] (Newton iteration on P_n).
void gauss_legendre(int n, std::vector<double>& x, std::vector<double>& w) {
  x.assign(n, 0.0);
  w.assign(n, 0.0);
  for (int i = 0; i < (n + 1) / 2; ++i) {
    double z = std::cos(M_PI * (i + 0.75) / (n + 0.5)), pp = 1.0;
    for (int it = 0; it < 100; ++it) {
      double p1 = 1.0, p2 = 0.0;
      for (int j = 1; j <= n; ++j) {
        double p3 = p2;
        p2 = p1;
        p1 = ((2.0 * j - 1.0) * z * p2 - (j - 1.0) * p3) / j;
      }
      pp = n * (z * p1 - p2) / (z * z - 1.0);
      double dz = p1 / pp;
      z -= dz;
      if (std::fabs(dz) < 1e-15) break;
    }
    x[i] = -z;
    x[n - 1 - i] = z;
    w[i] = w[n - 1 - i] = 2.0 / ((1.0 - z * z) * pp * pp);
  }
}

int correction_quadrature_size(int W) { return 2 * W + 40; }

// t = (W/2) sin(theta) removes the square-root end-point singularity of phi.
double kernel_ft_one(double xi, int W, double beta, const std::vector<double>& glx,
                     const std::vector<double>& glw) {
  const double bW = beta * W, h = 0.5 * W;
  double acc = 0.0;
  for (size_t q = 0; q < glx.size(); ++q) {
    double th = glx[q] * (M_PI / 2), ct = std::cos(th);
    acc += glw[q] * std::exp(bW * (ct - 1.0)) * ct * std::cos(2.0 * M_PI * h * std::sin(th) * xi);
  }
  return acc * h * (M_PI / 2);
}

GridGeom make_geom(const cip_plan& p) {
  GridGeom g{};
  g.px = p.pixsize_x;
  g.py = p.pixsize_y;
  g.w0 = p.w0;
  g.inv_dw = 1.0 / p.dw;
  g.nshift = p.nshift;
  g.nu = p.nu;
  g.nv = p.nv;
  g.W = p.support;
  g.nplanes = p.nplanes;
  g.do_wstacking = p.do_wstacking;
  g.beta_w = (float)(p.beta * p.support);
  KernelEntry ke;
  g.bf16_operands = kernel_lookup(p.sigma, p.support, &ke) && ke.err >= 5e-6;
  g.beta_wd = p.beta * p.support;
  if (p.support <= kFastSupport) {
    g.win_step = kWin - p.support + 1;
    int np0 = std::max(p.nplanes - p.support + 1, 1);  // number of distinct first planes
    g.nwin = (np0 + g.win_step - 1) / g.win_step;
  } else {
    g.win_step = 1;
    g.nwin = 1;
  }
  g.ntu = (p.nu + 2 * kHalo) / kTile + 1;
  g.ntv = (p.nv + 2 * kHalo) / kTile + 1;
  // origins i0r in 0..7 snapped to a lattice of 9 - W cells: ceil(8 / snap) positions per tile edge
  g.nlat = p.support <= kFastSupport ? (kTile + (9 - p.support) - 1) / (9 - p.support) : 1;
  g.win_begin = 0;
  g.win_end = g.nwin;
  return g;
}

}  // namespace cip

using namespace cip;

extern "C" cip_status cip_kernel_table(double sigma, int support, double* beta, double* rms_err) {
  KernelEntry e;
  CIP_REQUIRE(kernel_lookup(sigma, support, &e), "no kernel for sigma=%g W=%d", sigma, support);
  if (beta) *beta = e.beta;
  if (rms_err) *rms_err = e.err;
  return CIP_OK;
}

extern "C" cip_status cip_plan_make(int npix_x, int npix_y, double pixsize_x, double pixsize_y, double epsilon,
                                    int do_wstacking, int accumulate_fp64, double wmin, double wmax,
                                    int64_t nactive, double sigma_override, int support_override,
                                    cip_plan* out) {
  CIP_REQUIRE(out != nullptr, "plan output is NULL");
  CIP_REQUIRE(npix_x >= 16 && npix_y >= 16 && npix_x % 2 == 0 && npix_y % 2 == 0,
              "image dimensions must be even and >= 16 (got %d x %d)", npix_x, npix_y);
  CIP_REQUIRE(pixsize_x > 0 && pixsize_y > 0, "pixel sizes must be positive");
  CIP_REQUIRE(epsilon > 0 && epsilon < 1, "epsilon must be in (0,1)");
  CIP_REQUIRE(wmax >= wmin || nactive == 0, "wmax < wmin");
  const bool fp64 = accumulate_fp64 != 0, wst = do_wstacking != 0;
  const double eps1 = wst ? epsilon / (fp64 ? 30.0 : 3.0) : epsilon / (fp64 ? 20.0 : 2.0);
  double nm1_min = 0.0;
  if (wst) {
    double l = (npix_x / 2) * pixsize_x, m = (npix_y / 2) * pixsize_y, e = l * l + m * m;
    nm1_min = e >= 1.0 ? -1.0 : -e / (std::sqrt(1.0 - e) + 1.0);
  }
  const double nshift = -0.5 * nm1_min, half = 0.5 * std::fabs(nm1_min);
  const double scale = fp64 ? 2.0 : 1.0;
  bool found = false;
  double best_cost = 0;
  cip_plan best{};
  for (int s = 0; s < kNumSigma; ++s) {
    const double sg = kTable[s].sigma;
    if (sigma_override > 0 && std::fabs(sg - sigma_override) > 1e-9) continue;
    if (sigma_override <= 0 && !fp64 && sg < 1.4) continue;  // fp32 grids: keep 1/c(xi) edge gain modest
    int W = support_override;
    if (W <= 0) {
      for (int k = kMinSupport; k <= kMaxSupport; ++k)
        if (kTable[s].e[k - kMinSupport].err <= eps1) { W = k; break; }
      if (W <= 0) continue;
    }
    KernelEntry ke;
    CIP_REQUIRE(kernel_lookup(sg, W, &ke), "no kernel for sigma=%g W=%d", sg, W);
    int nu = good_size((int)std::ceil(npix_x * sg)), nv = good_size((int)std::ceil(npix_y * sg));
    double dw = 1.0;
    int npl = 1;
    if (wst && half != 0.0) {
      dw = 1.0 / (2.0 * sg * half);
      npl = (int)std::floor((wmax - wmin) / dw + W);
    }
    double per_vis = W <= kFastSupport ? kCostVisFast : (double)W * W * (wst ? W : 1) * kCostCellUpdate;
    double cost = scale * ((double)nactive * per_vis + (double)npl * nu * nv * kCostGridCell);
    if (!found || cost < best_cost) {
      found = true;
      best_cost = cost;
      best.sigma = sg;
      best.support = W;
      best.beta = ke.beta;
      best.nu = nu;
      best.nv = nv;
      best.dw = dw;
      best.nplanes = npl;
    }
  }
  CIP_REQUIRE(found, "epsilon=%g is too small for the kernel table", epsilon);
  best.npix_x = npix_x;
  best.npix_y = npix_y;
  best.pixsize_x = pixsize_x;
  best.pixsize_y = pixsize_y;
  best.nshift = nshift;
  best.w0 = (wst && half != 0.0) ? 0.5 * (wmin + wmax) - 0.5 * (best.nplanes - 1) * best.dw : 0.0;
  best.accumulate_fp64 = fp64;
  best.do_wstacking = wst;
  best.plane_block = 0;
  *out = best;
  return CIP_OK;
}

extern "C" cip_status cip_kernel_ft(const cip_plan* plan, const double* xi, int64_t n, double* out) {
  CIP_REQUIRE(plan && xi && out, "NULL argument");
  std::vector<double> x, w;
  gauss_legendre(correction_quadrature_size(plan->support), x, w);
  for (int64_t k = 0; k < n; ++k) out[k] = kernel_ft_one(xi[k], plan->support, plan->beta, x, w);
  return CIP_OK;
}
