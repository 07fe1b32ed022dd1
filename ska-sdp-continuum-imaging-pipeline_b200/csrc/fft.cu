// Kernel 3: per-w-plane 2-D FFT (stage G3), issued from C through cached cuFFT plans.
// Unnormalised, e^{+i...} sign (CUFFT_INVERSE), in place, one plan per (nu, nv, precision).
#include <cstdlib>

#include "common.h"

namespace cip {

namespace {

cip_status get_plan(cip_ctx* ctx, std::tuple<int, int, int, int> key, int rank, int* n, int* embed, int stride, int dist,
                    int batch, int fp64, cufftHandle* out) {
  auto it = ctx->fft_plans.find(key);
  if (it == ctx->fft_plans.end()) {
    FftPlan p;
    CIP_CHECK_CUFFT(cufftCreate(&p.handle));
    CIP_CHECK_CUFFT(cufftMakePlanMany(p.handle, rank, n, embed, stride, dist, embed, stride, dist,
                                      fp64 ? CUFFT_Z2Z : CUFFT_C2C, batch, &p.work));
    CIP_CHECK_CUFFT(cufftSetStream(p.handle, ctx->stream));
    it = ctx->fft_plans.emplace(key, p).first;
  }
  *out = it->second.handle;
  return CIP_OK;
}

cip_status exec(cufftHandle h, void* p, int fp64) {
  if (fp64) CIP_CHECK_CUFFT(cufftExecZ2Z(h, (cufftDoubleComplex*)p, (cufftDoubleComplex*)p, CUFFT_INVERSE));
  else CIP_CHECK_CUFFT(cufftExecC2C(h, (cufftComplex*)p, (cufftComplex*)p, CUFFT_INVERSE));
  return CIP_OK;
}

}  // namespace

// Contiguous batched 1-D transform (`batch` rows of length n), cached per (n, batch, precision).
cip_status fft_rows(cip_ctx* ctx, void* data, int n, int batch, int fp64) {
  int nn[1] = {n};
  cufftHandle h;
  CIP_TRY(get_plan(ctx, std::make_tuple(n, batch, -1, fp64), 1, nn, nn, 1, n, batch, fp64, &h));
  if (data) {
    CIP_TRY(exec(h, data, fp64));
    ctx->launches++;
  }
  return CIP_OK;
}

// keep_v: number of v columns of the transformed plane that the w-screen reads (the image width ny; they
// are the columns [0, ny/2) and [nv - ny/2, nv) in FFT order).  When keep_v < nv the transform is pruned:
// contiguous 1-D transforms along v for every row, then strided 1-D transforms along u for the retained
// columns only -- (nv - ny) / nv of the second pass is skipped.  Measured on B200 (6144^2 planes, 4096
// retained columns): 0.53 ms per plane against 0.46 ms for cuFFT's 2-D plan, whose strided pass is better
// than the strided 1-D plan -- so the pruned path is opt-in (CIP_FFT_PRUNED=1) and the default is the 2-D plan.
cip_status run_fft(cip_ctx* ctx, void* grid, int nu, int nv, int batch, int fp64, int keep_v) {
  const size_t csize = fp64 ? 16 : 8;
  const size_t plane = (size_t)nu * nv * csize;
  const bool pruned = keep_v > 0 && keep_v < nv && keep_v % 2 == 0 && getenv("CIP_FFT_PRUNED");
  if (!pruned) {
    int n[2] = {nu, nv};
    cufftHandle h;
    CIP_TRY(get_plan(ctx, std::make_tuple(nu, nv, 1, fp64), 2, n, nullptr, 1, nu * nv, 1, fp64, &h));
    ctx->launches += batch;
    for (int b = 0; b < batch; ++b) CIP_TRY(exec(h, static_cast<char*>(grid) + plane * b, fp64));
    return CIP_OK;
  }
  const int half = keep_v / 2;
  int nvv[1] = {nv}, nuu[1] = {nu};
  cufftHandle hv, hu;
  CIP_TRY(get_plan(ctx, std::make_tuple(nu, nv, 2, fp64), 1, nvv, nvv, 1, nv, nu, fp64, &hv));             // rows of one plane
  CIP_TRY(get_plan(ctx, std::make_tuple(nu, nv, 3 + half, fp64), 1, nuu, nuu, nv, 1, half, fp64, &hu));  // `half` columns
  ctx->launches += 3 * batch;
  for (int b = 0; b < batch; ++b) {
    char* p = static_cast<char*>(grid) + plane * b;
    CIP_TRY(exec(hv, p, fp64));
    CIP_TRY(exec(hu, p, fp64));
    CIP_TRY(exec(hu, p + (size_t)(nv - half) * csize, fp64));
  }
  return CIP_OK;
}

}  // namespace cip
