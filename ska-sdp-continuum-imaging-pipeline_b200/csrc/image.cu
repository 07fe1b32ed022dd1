// Kernel 4: fused w-screen / crop / accumulate per block of w planes, and the final grid
// correction (stage G4 of SURVEY.md section 8a; inside ducc0.wgridder.ms2dirty in the reference,
// call site src/ska_sdp_cip/invert.py:170-183; the scale factor folds in invert.py:149).
//
//   per plane p:  D[i,j] += Re( F_p[(i-nx/2) mod nu, (j-ny/2) mod nv] * exp(-2 pi i w_p (n-1+s)) )
//   once:         D[i,j] *= scale / ( c(x_i/nu) c(y_j/nv) c((n-1+s) dw) n )
//
// Both are HBM-bound streaming kernels: the w-screen pass reads only the nx*ny retained cells of
// each transformed plane and touches the image once per block of planes.
#include <algorithm>
#include <cmath>

#include "common.h"

namespace cip {

namespace {

struct ScreenParams {
  int nx, ny, nu, nv;
  double px, py, w0, dw, nshift;
  int plane_base, nslots, nplanes, do_wstacking;
};

template <typename R>
__global__ void __launch_bounds__(256) wscreen_kernel(ScreenParams p, const R* __restrict__ grid, R* __restrict__ image) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (j >= p.ny) return;
  const int gi = (i - p.nx / 2 + p.nu) % p.nu, gj = (j - p.ny / 2 + p.nv) % p.nv;
  const double l = (i - p.nx / 2) * p.px, m = (j - p.ny / 2) * p.py, e = l * l + m * m;
  if (e >= 1.0) return;
  const double z = p.do_wstacking ? -e / (sqrt(1.0 - e) + 1.0) + p.nshift : 0.0;
  const size_t plane = (size_t)p.nu * p.nv;
  const R* f = grid + 2 * ((size_t)gi * p.nv + gj);
  R acc = 0;
  for (int s = 0; s < p.nslots; ++s) {
    const int pl = p.plane_base + s;
    if (pl >= p.nplanes) break;
    double ph = (p.w0 + pl * p.dw) * z;
    ph -= rint(ph);
    R fr = f[2 * plane * s], fi = f[2 * plane * s + 1];
    if constexpr (sizeof(R) == 4) {
      float sn, cs;
      sincospif(2.0f * (float)ph, &sn, &cs);
      acc += fr * cs + fi * sn;
    } else {
      double sn, cs;
      sincospi(2.0 * ph, &sn, &cs);
      acc += fr * cs + fi * sn;
    }
  }
  image[(size_t)i * p.ny + j] += acc;
}

struct CorrectParams {
  int nx, ny;
  double px, py, dw, nshift, scale;
  double inv_xi2;  // 1 / ximax^2 of the Chebyshev fit
  int ncheb, do_wstacking, row_begin, row_end;
};

// tables: cu[nx], cv[ny], then ncheb Chebyshev coefficients of c(xi) in t = 2 (xi/ximax)^2 - 1
// (c is even and entire, so the series in xi^2 converges geometrically; Clenshaw recurrence).
template <typename R>
__global__ void __launch_bounds__(256) correct_kernel(CorrectParams p, const double* __restrict__ tables,
                                                      R* __restrict__ image) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = p.row_begin + blockIdx.y;
  if (j >= p.ny || i >= p.row_end) return;
  const double* cu = tables;
  const double* cv = tables + p.nx;
  const double* ck = cv + p.ny;
  const double l = (i - p.nx / 2) * p.px, m = (j - p.ny / 2) * p.py, e = l * l + m * m;
  const size_t k = (size_t)i * p.ny + j;
  if (e >= 1.0) { image[k] = 0; return; }
  double fac = cu[i] * cv[j];
  if (p.do_wstacking) {
    const double nm1 = -e / (sqrt(1.0 - e) + 1.0);
    const double xi = (nm1 + p.nshift) * p.dw;
    const double t = fmin(2.0 * xi * xi * p.inv_xi2 - 1.0, 1.0), t2 = 2.0 * t;
    double b1 = 0.0, b2 = 0.0;
    for (int q = p.ncheb - 1; q >= 1; --q) {
      double b0 = fma(t2, b1, ck[q] - b2);
      b2 = b1;
      b1 = b0;
    }
    const double cw = fma(t, b1, ck[0] - b2);
    fac *= cw * (nm1 + 1.0);
  }
  image[k] = (R)((double)image[k] * (p.scale / fac));
}

}  // namespace

cip_status run_wscreen(cip_ctx* ctx, const cip_plan& pl, const void* grid, int plane_base, int nslots, void* image) {
  ScreenParams p{pl.npix_x, pl.npix_y, pl.nu, pl.nv, pl.pixsize_x, pl.pixsize_y, pl.w0, pl.dw, pl.nshift,
                 plane_base, nslots, pl.nplanes, pl.do_wstacking};
  dim3 grd((pl.npix_y + 255) / 256, pl.npix_x);
  if (pl.accumulate_fp64)
    wscreen_kernel<double><<<grd, 256, 0, ctx->stream>>>(p, static_cast<const double*>(grid), static_cast<double*>(image));
  else
    wscreen_kernel<float><<<grd, 256, 0, ctx->stream>>>(p, static_cast<const float*>(grid), static_cast<float*>(image));
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

// Chebyshev coefficients of c(xi) on |xi| <= ximax in the variable t = 2 (xi/ximax)^2 - 1.
static std::vector<double> kernel_ft_chebyshev(int W, double beta, double ximax) {
  std::vector<double> glx, glw;
  gauss_legendre(correction_quadrature_size(W), glx, glw);
  constexpr int N = 64;
  double f[N];
  for (int k = 0; k < N; ++k) {
    double t = std::cos(M_PI * (k + 0.5) / N);
    f[k] = kernel_ft_one(ximax * std::sqrt(0.5 * (t + 1.0)), W, beta, glx, glw);
  }
  std::vector<double> c(N);
  for (int q = 0; q < N; ++q) {
    double a = 0.0;
    for (int k = 0; k < N; ++k) a += f[k] * std::cos(M_PI * q * (k + 0.5) / N);
    c[q] = a * (q == 0 ? 1.0 : 2.0) / N;
  }
  int n = N;
  while (n > 2 && std::fabs(c[n - 1]) < 1e-17 * std::fabs(c[0])) --n;
  c.resize(n);
  return c;
}

static double clenshaw(const std::vector<double>& c, double t) {
  double b1 = 0.0, b2 = 0.0;
  for (int q = (int)c.size() - 1; q >= 1; --q) {
    double b0 = 2.0 * t * b1 - b2 + c[q];
    b2 = b1;
    b1 = b0;
  }
  return t * b1 - b2 + c[0];
}

cip_status run_correct(cip_ctx* ctx, const cip_plan& pl, void* image, int row_begin, int row_end, double scale) {
  CIP_REQUIRE(row_begin >= 0 && row_end <= pl.npix_x && row_begin <= row_end, "bad row range [%d,%d)", row_begin, row_end);
  if (row_begin == row_end) return CIP_OK;
  const int nx = pl.npix_x, ny = pl.npix_y, W = pl.support;
  // one fit of c(xi) serves the u, v and w axes: |x_i/nu| <= nx/(2 nu), |(n-1+s) dw| <= 1/(2 sigma)
  double nm1_min = 0.0;
  {
    double l = (nx / 2) * pl.pixsize_x, m = (ny / 2) * pl.pixsize_y, e = l * l + m * m;
    nm1_min = e >= 1.0 ? -1.0 : -e / (std::sqrt(1.0 - e) + 1.0);
  }
  const double xi_w = pl.do_wstacking ? std::max(std::fabs(nm1_min + pl.nshift), std::fabs(pl.nshift)) * pl.dw : 0.0;
  const double ximax = 1.0001 * std::max({0.5 * nx / pl.nu, 0.5 * ny / pl.nv, xi_w, 1e-3});
  const std::vector<double> key = {(double)nx, (double)ny, (double)pl.nu, (double)pl.nv, (double)W, pl.beta, ximax};
  if (key != ctx->tables_key) {
    const std::vector<double> ck = kernel_ft_chebyshev(W, pl.beta, ximax);
    std::vector<double> tab((size_t)nx + ny + ck.size());
    auto c_of = [&](double xi) { return clenshaw(ck, 2.0 * (xi / ximax) * (xi / ximax) - 1.0); };
    for (int i = 0; i < nx; ++i) tab[i] = c_of((double)(i - nx / 2) / pl.nu);
    for (int j = 0; j < ny; ++j) tab[nx + j] = c_of((double)(j - ny / 2) / pl.nv);
    for (size_t q = 0; q < ck.size(); ++q) tab[nx + ny + q] = ck[q];
    CIP_TRY(ctx->tables.ensure(sizeof(double) * tab.size()));
    CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->tables.p, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice, ctx->stream));
    CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));  // tab is a local
    ctx->tables_key = key;
    ctx->tables_ncheb = (int)ck.size();
  }
  CorrectParams p{nx, ny, pl.pixsize_x, pl.pixsize_y, pl.dw, pl.nshift, scale, 1.0 / (ximax * ximax),
                  ctx->tables_ncheb, pl.do_wstacking, row_begin, row_end};
  dim3 grd((ny + 255) / 256, row_end - row_begin);
  if (pl.accumulate_fp64)
    correct_kernel<double><<<grd, 256, 0, ctx->stream>>>(p, ctx->tables.as<double>(), static_cast<double*>(image));
  else
    correct_kernel<float><<<grd, 256, 0, ctx->stream>>>(p, ctx->tables.as<double>(), static_cast<float*>(image));
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

}  // namespace cip
