// Kernel 3: per-w-plane 2-D FFT (stage G3), issued from C through cached cuFFT plans.
// Unnormalised, e^{+i...} sign (CUFFT_INVERSE), in place, one plan per (nu, nv, precision).
#include "common.h"

namespace cip {

cip_status run_fft(cip_ctx* ctx, void* grid, int nu, int nv, int batch, int fp64) {
  auto key = std::make_tuple(nu, nv, 1, fp64);
  auto it = ctx->fft_plans.find(key);
  if (it == ctx->fft_plans.end()) {
    FftPlan p;
    CIP_CHECK_CUFFT(cufftCreate(&p.handle));
    int n[2] = {nu, nv};
    CIP_CHECK_CUFFT(cufftMakePlanMany(p.handle, 2, n, nullptr, 1, nu * nv, nullptr, 1, nu * nv,
                                      fp64 ? CUFFT_Z2Z : CUFFT_C2C, 1, &p.work));
    CIP_CHECK_CUFFT(cufftSetStream(p.handle, ctx->stream));
    it = ctx->fft_plans.emplace(key, p).first;
  }
  ctx->launches += batch;
  const size_t plane = (size_t)nu * nv * (fp64 ? 16 : 8);
  for (int b = 0; b < batch; ++b) {
    void* p = static_cast<char*>(grid) + plane * b;
    if (fp64)
      CIP_CHECK_CUFFT(cufftExecZ2Z(it->second.handle, (cufftDoubleComplex*)p, (cufftDoubleComplex*)p, CUFFT_INVERSE));
    else
      CIP_CHECK_CUFFT(cufftExecC2C(it->second.handle, (cufftComplex*)p, (cufftComplex*)p, CUFFT_INVERSE));
  }
  return CIP_OK;
}

}  // namespace cip
