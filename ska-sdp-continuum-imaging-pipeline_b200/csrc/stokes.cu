// Stokes-I prologue (P1/P2): replaces the host numpy passes of
// StokesIGridderInput.from_measurement_set_reader and effective_weights
// (src/ska_sdp_cip/invert.py:72-116) with one streaming pass on the device:
//   vis_I = 0.5 (v0 + v3);  eff_w = !(f0 | f3) * 4 / (1/w0 + 1/w3)
// float32 arithmetic in numpy's operation order (1/w0 and 1/w3 are inf for zero weights, so
// eff_w is 0), HBM-bound: 52 B in, 12 B out per sample.
#include <algorithm>

#include "common.h"

namespace cip {

__global__ void __launch_bounds__(256) stokes_kernel(const float4* __restrict__ vis4 /* 2 float4 per sample */,
                                                     const uint32_t* __restrict__ flags4 /* 4 bytes per sample */,
                                                     const float4* __restrict__ wgt4, int64_t n,
                                                     float2* __restrict__ vis_out, float* __restrict__ effw_out) {
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
    float4 a = vis4[2 * k], b = vis4[2 * k + 1];  // (v0, v1), (v2, v3)
    float2 v;
    v.x = __fmul_rn(0.5f, __fadd_rn(a.x, b.z));
    v.y = __fmul_rn(0.5f, __fadd_rn(a.y, b.w));
    vis_out[k] = v;
    uint32_t f = flags4[k];
    bool flagged = (f & 0x000000ffu) || (f & 0xff000000u);
    float4 w = wgt4[k];
    float s = __fadd_rn(__fdiv_rn(1.0f, w.x), __fdiv_rn(1.0f, w.w));
    float wi = __fdiv_rn(4.0f, s);
    effw_out[k] = flagged ? 0.0f : wi;
  }
}

}  // namespace cip

using namespace cip;

extern "C" cip_status cip_stokes_i(cip_ctx* ctx, const void* vis4, const uint8_t* flags4, const float* wgt4,
                                   int64_t nrow, int64_t nchan, void* vis_out, float* effw_out, int memspace) {
  CIP_REQUIRE(ctx && vis4 && flags4 && wgt4 && vis_out && effw_out, "NULL argument");
  CIP_REQUIRE(nrow >= 0 && nchan >= 0, "negative nrow/nchan");
  CIP_REQUIRE(memspace == 0 || memspace == 1, "memspace must be 0 or 1");
  CIP_ENTER_DEVICE(ctx);
  const int64_t n = nrow * nchan;
  if (n == 0) return CIP_OK;
  cudaStream_t st = ctx->stream;
  const void* dv = vis4;
  const void* df = flags4;
  const void* dw = wgt4;
  void* ov = vis_out;
  void* ow = effw_out;
  if (memspace == 0) {
    CIP_TRY(ctx->t_vis.ensure((size_t)n * (32 + 4 + 16)));
    char* base = ctx->t_vis.as<char>();
    CIP_CHECK_CUDA(cudaMemcpyAsync(base, vis4, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    CIP_CHECK_CUDA(cudaMemcpyAsync(base + (size_t)n * 32, wgt4, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    CIP_CHECK_CUDA(cudaMemcpyAsync(base + (size_t)n * 48, flags4, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    dv = base;
    dw = base + (size_t)n * 32;
    df = base + (size_t)n * 48;
    ctx->staged_valid = false;  // the outputs overwrite the staged inputs of a preceding cip_scan
    CIP_TRY(ctx->in_vis.ensure((size_t)n * 8));
    CIP_TRY(ctx->in_wgt.ensure((size_t)n * 4));
    ov = ctx->in_vis.p;
    ow = ctx->in_wgt.p;
  }
  int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
  stokes_kernel<<<blocks, 256, 0, st>>>(static_cast<const float4*>(dv), static_cast<const uint32_t*>(df),
                                        static_cast<const float4*>(dw), n, static_cast<float2*>(ov),
                                        static_cast<float*>(ow));
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  if (memspace == 0) {
    CIP_CHECK_CUDA(cudaMemcpyAsync(vis_out, ov, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    CIP_CHECK_CUDA(cudaMemcpyAsync(effw_out, ow, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  }
  CIP_CHECK_CUDA(cudaStreamSynchronize(st));
  return CIP_OK;
}

// The same on device-resident blocks without waiting for the kernel: the caller orders it against its uploads with
// events on cip_ctx_stream (the pinned double-buffered reader pipeline of ska_sdp_cip.invert).
extern "C" cip_status cip_stokes_i_async(cip_ctx* ctx, const void* vis4, const uint8_t* flags4, const float* wgt4, int64_t nrow,
                                         int64_t nchan, void* vis_out, float* effw_out) {
  CIP_REQUIRE(ctx && vis4 && flags4 && wgt4 && vis_out && effw_out, "NULL argument");
  CIP_REQUIRE(nrow >= 0 && nchan >= 0, "negative nrow/nchan");
  CIP_ENTER_DEVICE(ctx);
  const int64_t n = nrow * nchan;
  if (n == 0) return CIP_OK;
  int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
  stokes_kernel<<<blocks, 256, 0, ctx->stream>>>(static_cast<const float4*>(vis4), reinterpret_cast<const uint32_t*>(flags4),
                                                 reinterpret_cast<const float4*>(wgt4), n, static_cast<float2*>(vis_out), effw_out);
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}
