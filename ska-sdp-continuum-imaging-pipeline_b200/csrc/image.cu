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
#include <cstdlib>

#include "common.h"

namespace cip {

namespace {

struct ScreenParams {
  int nx, ny, nu, nv;
  double px, py, w0, dw, nshift;
  int plane_base, nslots, nplanes, do_wstacking;
  const uint32_t* skip;  // != 0 on the device: a binning pass of this call overflowed, the image must stay untouched
};

template <typename R>
__global__ void __launch_bounds__(256) wscreen_kernel(ScreenParams p, const R* __restrict__ grid, R* __restrict__ image) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (j >= p.ny || *p.skip) return;
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

template <typename R> struct Cplx;
template <> struct Cplx<float> { using type = float2; };
template <> struct Cplx<double> { using type = double2; };

// T[j][iu] = F[iu][(j - ny/2) mod nv]: the ny retained v columns of a plane that has been transformed along v,
// transposed so that the transforms along u are contiguous rows.  32 x 32 tiles through shared memory.
template <typename R>
__global__ void __launch_bounds__(256) transpose_crop_kernel(const typename Cplx<R>::type* __restrict__ f,
                                                             typename Cplx<R>::type* __restrict__ t, int nu, int nv, int ny) {
  using C = typename Cplx<R>::type;
  __shared__ C tile[32][33];
  const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int i = i0 + r, j = j0 + tx;
    if (i < nu && j < ny) {
      int gj = j - ny / 2;
      gj += gj < 0 ? nv : 0;
      tile[r][tx] = f[(size_t)i * nv + gj];
    }
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int j = j0 + r, i = i0 + tx;
    if (i < nu && j < ny) t[(size_t)j * nu + i] = tile[tx][r];
  }
}

// D[i,j] += Re( T[j][(i - nx/2) mod nu] * exp(-2 pi i w_p (n-1+s)) ) for one transformed, transposed plane.
template <typename R>
__global__ void __launch_bounds__(256) wscreen_t_kernel(ScreenParams p, int plane, const typename Cplx<R>::type* __restrict__ t,
                                                        R* __restrict__ image) {
  __shared__ R tile[32][33];
  if (*p.skip) return;
  const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const double wp = p.w0 + plane * p.dw;
  for (int r = ty; r < 32; r += 8) {  // lanes along i: coalesced reads of T
    const int j = j0 + r, i = i0 + tx;
    R val = 0;
    if (i < p.nx && j < p.ny) {
      const double l = (i - p.nx / 2) * p.px, m = (j - p.ny / 2) * p.py, e = l * l + m * m;
      if (e < 1.0) {
        int gi = i - p.nx / 2;
        gi += gi < 0 ? p.nu : 0;
        const auto f = t[(size_t)j * p.nu + gi];
        const double z = p.do_wstacking ? -e / (sqrt(1.0 - e) + 1.0) + p.nshift : 0.0;
        double ph = wp * z;
        ph -= rint(ph);
        if constexpr (sizeof(R) == 4) {
          float sn, cs;
          sincospif(2.0f * (float)ph, &sn, &cs);
          val = f.x * cs + f.y * sn;
        } else {
          double sn, cs;
          sincospi(2.0 * ph, &sn, &cs);
          val = f.x * cs + f.y * sn;
        }
      }
    }
    tile[r][tx] = val;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {  // lanes along j: coalesced image update
    const int i = i0 + r, j = j0 + tx;
    if (i < p.nx && j < p.ny) image[(size_t)i * p.ny + j] += tile[tx][r];
  }
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
  // the Chebyshev coefficients are read by every thread in the same order: shared memory (one broadcast
  // load each) instead of ncheb global loads per pixel, which bound the kernel
  __shared__ double ck[64];
  for (int q = threadIdx.x; q < p.ncheb; q += blockDim.x) ck[q] = tables[p.nx + p.ny + q];
  __syncthreads();
  // the factor is even in m: one thread serves the columns j and ny - j (bit-identical to one thread per pixel)
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = p.row_begin + blockIdx.y;
  if (j > p.ny / 2 || i >= p.row_end) return;
  const double* cu = tables;
  const double* cv = tables + p.nx;
  const double l = (i - p.nx / 2) * p.px, m = (j - p.ny / 2) * p.py, e = l * l + m * m;
  const size_t k = (size_t)i * p.ny + j;
  const bool twin = j > 0 && 2 * j < p.ny;
  const size_t k2 = (size_t)i * p.ny + (p.ny - j);
  if (e >= 1.0) {
    image[k] = 0;
    if (twin) image[k2] = 0;
    return;
  }
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
  const double f = p.scale / fac;
  image[k] = (R)((double)image[k] * f);
  if (twin) image[k2] = (R)((double)image[k2] * f);
}

}  // namespace

cip_status run_wscreen(cip_ctx* ctx, const cip_plan& pl, const void* grid, int plane_base, int nslots, void* image) {
  ScreenParams p{pl.npix_x, pl.npix_y, pl.nu, pl.nv, pl.pixsize_x, pl.pixsize_y, pl.w0, pl.dw, pl.nshift,
                 plane_base, nslots, pl.nplanes, pl.do_wstacking, ctx->call_flags.as<uint32_t>()};
  dim3 grd((pl.npix_y + 255) / 256, pl.npix_x);
  if (pl.accumulate_fp64)
    wscreen_kernel<double><<<grd, 256, 0, ctx->stream>>>(p, static_cast<const double*>(grid), static_cast<double*>(image));
  else
    wscreen_kernel<float><<<grd, 256, 0, ctx->stream>>>(p, static_cast<const float*>(grid), static_cast<float*>(image));
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

template <typename R>
static cip_status fft_screen_transposed(cip_ctx* ctx, const cip_plan& pl, const ScreenParams& p, void* grid, int nslots,
                                        void* image) {
  using C = typename Cplx<R>::type;
  const int nu = pl.nu, nv = pl.nv, nx = pl.npix_x, ny = pl.npix_y;
  CIP_TRY(ctx->fft_t.ensure(sizeof(C) * (size_t)ny * nu));
  C* t = ctx->fft_t.as<C>();
  const dim3 tg((ny + 31) / 32, (nu + 31) / 32), sg((ny + 31) / 32, (nx + 31) / 32);
  for (int s = 0; s < nslots; ++s) {
    const int plane = p.plane_base + s;
    if (plane >= p.nplanes) break;
    C* f = static_cast<C*>(grid) + (size_t)s * nu * nv;
    CIP_TRY(fft_rows(ctx, f, nv, nu, sizeof(R) == 8));                     // along v, every row
    transpose_crop_kernel<R><<<tg, 256, 0, ctx->stream>>>(f, t, nu, nv, ny);  // retained columns only
    CIP_TRY(fft_rows(ctx, t, nu, ny, sizeof(R) == 8));                     // along u, ny rows of the transpose
    wscreen_t_kernel<R><<<sg, 256, 0, ctx->stream>>>(p, plane, t, static_cast<R*>(image));
    ctx->launches += 2;
  }
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

cip_status run_fft_screen(cip_ctx* ctx, const cip_plan& pl, void* grid, int plane_base, int nslots, void* image,
                          double* ms_fft, double* ms_wscreen) {
  NvtxRange nvtx("cip:fft+wscreen");
  (void)ms_wscreen;
  bool own = false;
  if (!getenv("CIP_FFT_2D")) CIP_TRY(pfft_prepare(ctx, pl, &own));  // the own pruned transform with the fused w-screen (pfft.cu)
  if (own) return grid ? run_pfft_screen(ctx, pl, grid, plane_base, nslots, image) : CIP_OK;
  if (grid == nullptr) {  // plans only
    if (getenv("CIP_FFT_2D")) return run_fft(ctx, nullptr, pl.nu, pl.nv, 0, pl.accumulate_fp64, pl.npix_y);
    // plans, their work areas and the transposed plane exist before the plane stack is sized from what is free
    CIP_TRY(ctx->fft_t.ensure((size_t)(pl.accumulate_fp64 ? 16 : 8) * pl.npix_y * pl.nu));
    CIP_TRY(fft_rows(ctx, nullptr, pl.nv, pl.nu, pl.accumulate_fp64));
    return fft_rows(ctx, nullptr, pl.nu, pl.npix_y, pl.accumulate_fp64);
  }
  (void)ms_fft;
  if (getenv("CIP_FFT_2D")) {
    CIP_TRY(run_fft(ctx, grid, pl.nu, pl.nv, std::min(nslots, pl.nplanes - plane_base), pl.accumulate_fp64, pl.npix_y));
    return run_wscreen(ctx, pl, grid, plane_base, nslots, image);
  }
  ScreenParams p{pl.npix_x, pl.npix_y, pl.nu, pl.nv, pl.pixsize_x, pl.pixsize_y, pl.w0, pl.dw, pl.nshift,
                 plane_base, nslots, pl.nplanes, pl.do_wstacking, ctx->call_flags.as<uint32_t>()};
  if (pl.accumulate_fp64) return fft_screen_transposed<double>(ctx, pl, p, grid, nslots, image);
  return fft_screen_transposed<float>(ctx, pl, p, grid, nslots, image);
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
  NvtxRange nvtx("cip:correction");
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
  dim3 grd((ny / 2 + 1 + 255) / 256, row_end - row_begin);
  if (pl.accumulate_fp64)
    correct_kernel<double><<<grd, 256, 0, ctx->stream>>>(p, ctx->tables.as<double>(), static_cast<double*>(image));
  else
    correct_kernel<float><<<grd, 256, 0, ctx->stream>>>(p, ctx->tables.as<double>(), static_cast<float*>(image));
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

}  // namespace cip
