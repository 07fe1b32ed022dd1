// Kernel 4: fused w-screen / crop / accumulate per block of w planes, and the final grid
// correction (stage G4 of SURVEY.md section 8a; inside ducc0.wgridder.ms2dirty in the reference,
// call site src/ska_sdp_cip/invert.py:170-183; the scale factor folds in invert.py:149).
//
//   per plane p:  D[i,j] += Re( F_p[(i-nx/2) mod nu, (j-ny/2) mod nv] * exp(-2 pi i w_p (n-1+s)) )
//   once:         D[i,j] *= scale / ( c(x_i/nu) c(y_j/nv) c((n-1+s) dw) n )
//
// Both are HBM-bound streaming kernels: the w-screen pass reads only the nx*ny retained cells of
// each transformed plane and touches the image once per block of planes.
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
  int nq, do_wstacking, row_begin, row_end;
};

// tables: cu[nx], cv[ny], then nq quadrature abscissae t_q and nq weights a_q with
// c(xi) = sum_q a_q cos(2 pi t_q xi)
template <typename R>
__global__ void __launch_bounds__(256) correct_kernel(CorrectParams p, const double* __restrict__ tables,
                                                      R* __restrict__ image) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = p.row_begin + blockIdx.y;
  if (j >= p.ny || i >= p.row_end) return;
  const double* cu = tables;
  const double* cv = tables + p.nx;
  const double* tq = cv + p.ny;
  const double* aq = tq + p.nq;
  const double l = (i - p.nx / 2) * p.px, m = (j - p.ny / 2) * p.py, e = l * l + m * m;
  const size_t k = (size_t)i * p.ny + j;
  if (e >= 1.0) { image[k] = 0; return; }
  double fac = cu[i] * cv[j];
  if (p.do_wstacking) {
    const double nm1 = -e / (sqrt(1.0 - e) + 1.0);
    const double xi = (nm1 + p.nshift) * p.dw;
    double cw = 0.0;
    for (int q = 0; q < p.nq; ++q) cw += aq[q] * cospi(2.0 * tq[q] * xi);
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

cip_status run_correct(cip_ctx* ctx, const cip_plan& pl, void* image, int row_begin, int row_end, double scale) {
  CIP_REQUIRE(row_begin >= 0 && row_end <= pl.npix_x && row_begin <= row_end, "bad row range [%d,%d)", row_begin, row_end);
  if (row_begin == row_end) return CIP_OK;
  const int nx = pl.npix_x, ny = pl.npix_y, W = pl.support;
  std::vector<double> glx, glw;
  gauss_legendre(correction_quadrature_size(W), glx, glw);
  // half of the symmetric nodes suffices: c(xi) = sum_{q: x_q > 0} 2 a_q cos(2 pi t_q xi)
  std::vector<double> tq, aq;
  const double bW = pl.beta * W, h = 0.5 * W;
  for (size_t q = 0; q < glx.size(); ++q) {
    if (glx[q] < 0.0) continue;
    double th = glx[q] * (M_PI / 2), ct = std::cos(th);
    double a = glw[q] * std::exp(bW * (ct - 1.0)) * ct * h * (M_PI / 2);
    tq.push_back(h * std::sin(th));
    aq.push_back(glx[q] == 0.0 ? a : 2.0 * a);
  }
  const int nq = (int)tq.size();
  std::vector<double> tab((size_t)nx + ny + 2 * nq);
  for (int i = 0; i < nx; ++i) tab[i] = kernel_ft_one((double)(i - nx / 2) / pl.nu, W, pl.beta, glx, glw);
  for (int j = 0; j < ny; ++j) tab[nx + j] = kernel_ft_one((double)(j - ny / 2) / pl.nv, W, pl.beta, glx, glw);
  for (int q = 0; q < nq; ++q) { tab[nx + ny + q] = tq[q]; tab[nx + ny + nq + q] = aq[q]; }
  CIP_TRY(ctx->tables.ensure(sizeof(double) * tab.size()));
  CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->tables.p, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice, ctx->stream));
  CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));  // tab is a local
  CorrectParams p{nx, ny, pl.pixsize_x, pl.pixsize_y, pl.dw, pl.nshift, scale, nq, pl.do_wstacking, row_begin, row_end};
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
