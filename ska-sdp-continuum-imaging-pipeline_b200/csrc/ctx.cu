// Context, error reporting, device buffers and input staging of libcip_b200.so.
#include <cstring>

#include "common.h"

namespace cip {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

cip_status DevBuf::ensure(size_t bytes) {
  if (bytes <= cap) return CIP_OK;
  release();
  size_t want = bytes + bytes / 16 + 256;  // a little slack so slowly growing shards do not thrash
  cudaError_t e = cudaMalloc(&p, want);
  if (e != cudaSuccess) {
    cudaGetLastError();
    want = bytes;
    e = cudaMalloc(&p, want);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    p = nullptr;
    cap = 0;
    set_error("cudaMalloc of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    return e == cudaErrorMemoryAllocation ? CIP_ENOMEM : CIP_ECUDA;
  }
  cap = want;
  return CIP_OK;
}

void DevBuf::release() {
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
}

float elapsed_ms(cip_ctx* ctx, int e0, int e1) {
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx->ev[e0], ctx->ev[e1]);
  return ms;
}

cip_status validate_args(const cip_invert_args* a) {
  CIP_REQUIRE(a != nullptr, "args is NULL");
  CIP_REQUIRE(a->nrow >= 0 && a->nchan >= 0, "negative nrow/nchan");
  CIP_REQUIRE(a->nrow < (int64_t)1 << 31 && a->nchan < (int64_t)1 << 20, "nrow/nchan too large");
  CIP_REQUIRE(a->nrow == 0 || a->nchan == 0 || (a->uvw && a->freq && a->vis), "uvw, freq and vis must not be NULL");
  CIP_REQUIRE(a->npix_x >= 16 && a->npix_y >= 16 && a->npix_x % 2 == 0 && a->npix_y % 2 == 0,
              "image dimensions must be even and >= 16 (got %d x %d)", a->npix_x, a->npix_y);
  CIP_REQUIRE(a->pixsize_x > 0 && a->pixsize_y > 0, "pixel sizes must be positive");
  CIP_REQUIRE(a->epsilon > 0 && a->epsilon < 1, "epsilon must be in (0,1)");
  CIP_REQUIRE(a->vis_row_stride == 0 || a->vis_row_stride >= a->nchan, "vis_row_stride < nchan");
  CIP_REQUIRE(a->wgt_row_stride == 0 || a->wgt_row_stride >= a->nchan, "wgt_row_stride < nchan");
  CIP_REQUIRE(a->memspace == 0 || a->memspace == 1, "memspace must be 0 (host) or 1 (device)");
  return CIP_OK;
}

// Host inputs are copied to context-owned device buffers (rows compacted to stride nchan);
// device inputs are used in place.
cip_status stage_inputs(cip_ctx* ctx, const cip_invert_args* a, DevInputs* d, double* ms_h2d) {
  const int64_t nrow = a->nrow, nchan = a->nchan;
  const int64_t vs = a->vis_row_stride ? a->vis_row_stride : nchan;
  const int64_t ws = a->wgt_row_stride ? a->wgt_row_stride : nchan;
  d->nrow = nrow;
  d->nchan = nchan;
  if (ms_h2d) *ms_h2d = 0.0;
  if (a->memspace == 1) {
    d->uvw = a->uvw;
    d->freq = a->freq;
    d->vis = static_cast<const float2*>(a->vis);
    d->wgt = a->wgt;
    d->mask = a->mask;
    d->vstride = vs;
    d->wstride = ws;
    return CIP_OK;
  }
  cudaStream_t st = ctx->stream;
  CIP_CHECK_CUDA(cudaEventRecord(ctx->ev[14], st));
  CIP_TRY(ctx->in_uvw.ensure(sizeof(double) * 3 * (size_t)nrow + 8));
  CIP_TRY(ctx->in_freq.ensure(sizeof(double) * (size_t)nchan + 8));
  CIP_TRY(ctx->in_vis.ensure(sizeof(float2) * (size_t)nrow * nchan + 8));
  if (nrow && nchan) {
    CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_uvw.p, a->uvw, sizeof(double) * 3 * nrow, cudaMemcpyHostToDevice, st));
    CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_freq.p, a->freq, sizeof(double) * nchan, cudaMemcpyHostToDevice, st));
    CIP_CHECK_CUDA(cudaMemcpy2DAsync(ctx->in_vis.p, sizeof(float2) * nchan, a->vis, sizeof(float2) * vs,
                                     sizeof(float2) * nchan, nrow, cudaMemcpyHostToDevice, st));
  }
  d->uvw = ctx->in_uvw.as<double>();
  d->freq = ctx->in_freq.as<double>();
  d->vis = ctx->in_vis.as<float2>();
  d->vstride = nchan;
  d->wstride = nchan;
  d->wgt = nullptr;
  d->mask = nullptr;
  if (a->wgt) {
    CIP_TRY(ctx->in_wgt.ensure(sizeof(float) * (size_t)nrow * nchan + 8));
    if (nrow && nchan)
      CIP_CHECK_CUDA(cudaMemcpy2DAsync(ctx->in_wgt.p, sizeof(float) * nchan, a->wgt, sizeof(float) * ws,
                                       sizeof(float) * nchan, nrow, cudaMemcpyHostToDevice, st));
    d->wgt = ctx->in_wgt.as<float>();
  }
  if (a->mask) {
    CIP_TRY(ctx->in_mask.ensure((size_t)nrow * nchan + 8));
    if (nrow && nchan)
      CIP_CHECK_CUDA(cudaMemcpy2DAsync(ctx->in_mask.p, nchan, a->mask, ws, nchan, nrow, cudaMemcpyHostToDevice, st));
    d->mask = ctx->in_mask.as<uint8_t>();
  }
  CIP_CHECK_CUDA(cudaEventRecord(ctx->ev[15], st));
  if (ms_h2d) {
    CIP_CHECK_CUDA(cudaEventSynchronize(ctx->ev[15]));
    *ms_h2d = elapsed_ms(ctx, 14, 15);
  }
  return CIP_OK;
}

}  // namespace cip

using namespace cip;

extern "C" const char* cip_last_error(void) { return g_err; }
extern "C" const char* cip_version(void) { return "cip_b200 0.1.0 (sm_100a)"; }

extern "C" cip_status cip_ctx_create(int device, cip_ctx** out) {
  CIP_REQUIRE(out != nullptr, "ctx output is NULL");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    set_error("no CUDA device available (%s); libcip_b200 has no CPU fallback",
              e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    return CIP_ECUDA;
  }
  if (device < 0) CIP_CHECK_CUDA(cudaGetDevice(&device));
  CIP_REQUIRE(device < ndev, "device %d out of range (%d devices)", device, ndev);
  CIP_CHECK_CUDA(cudaSetDevice(device));
  cip_ctx* ctx = new cip_ctx();
  ctx->device = device;
  cudaDeviceProp prop{};
  CIP_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
  ctx->sm_count = prop.multiProcessorCount;
  ctx->smem_optin = prop.sharedMemPerBlockOptin;
  CIP_CHECK_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  for (auto& ev : ctx->ev) CIP_CHECK_CUDA(cudaEventCreate(&ev));
  CIP_CHECK_CUDA(cudaMallocHost(&ctx->scan_host, 256));
  *out = ctx;
  return CIP_OK;
}

extern "C" cip_status cip_ctx_destroy(cip_ctx* ctx) {
  if (!ctx) return CIP_OK;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  for (auto& kv : ctx->fft_plans) cufftDestroy(kv.second.handle);
  DevBuf* bufs[] = {&ctx->in_uvw, &ctx->in_freq, &ctx->in_vis, &ctx->in_wgt, &ctx->in_mask, &ctx->out_img,
                    &ctx->scan_out, &ctx->row_nruns, &ctx->row_off, &ctx->runs, &ctx->run_keys,
                    &ctx->run_keys_alt, &ctx->run_idx, &ctx->run_idx_alt, &ctx->runs_sorted, &ctx->run_len,
                    &ctx->run_visoff, &ctx->flags, &ctx->group_start, &ctx->item_idx, &ctx->items,
                    &ctx->win_item_start, &ctx->cub_tmp, &ctx->recs, &ctx->grid, &ctx->image, &ctx->tables,
                    &ctx->t_uvw, &ctx->t_freq, &ctx->t_vis};
  for (DevBuf* b : bufs) b->release();
  if (ctx->scan_host) cudaFreeHost(ctx->scan_host);
  for (auto& ev : ctx->ev)
    if (ev) cudaEventDestroy(ev);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return CIP_OK;
}

extern "C" void* cip_ctx_stream(cip_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

extern "C" int64_t cip_ctx_launches(cip_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" cip_status cip_ctx_synchronize(cip_ctx* ctx) {
  CIP_REQUIRE(ctx != nullptr, "ctx is NULL");
  CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
  return CIP_OK;
}
