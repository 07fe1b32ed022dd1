// Context, error reporting, device buffers and input staging of libcip_b200.so.
#include <cstdlib>
#include <cstring>
#include <ctime>

#include <algorithm>

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

double now_ms() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

void trace_mark(const char* label) {
  static const bool on = getenv("CIP_TRACE") != nullptr;
  if (!on) return;
  static double t0 = 0;
  const double t = now_ms();
  if (t0 == 0 || strcmp(label, "begin") == 0) t0 = t;
  fprintf(stderr, "[cip trace] %8.3f ms  %s\n", t - t0, label);
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
  if (a->row_vis_offset || a->row_chan_start || a->row_chan_stop) {
    CIP_REQUIRE(a->row_vis_offset && a->row_chan_start && a->row_chan_stop, "jagged rows need all three row tables");
    CIP_REQUIRE(a->nvis >= 0 && a->nvis < (int64_t)1 << 32, "bad nvis for jagged rows");
  }
  CIP_REQUIRE(!(a->w_hi > 0) || (a->do_wstacking && a->w_lo >= 0 && a->w_lo < a->w_hi), "bad |w| slab [%g, %g)", a->w_lo, a->w_hi);
  return CIP_OK;
}

// nrow rows of nchan elements, source row stride `stride` elements -> packed rows on the device.  Packed
// sources go as ONE linear copy: a 2-D copy of millions of short rows is descriptor-bound.
static cip_status copy_rows(void* dst, const void* src, size_t elem, int64_t nrow, int64_t nchan, int64_t stride,
                            cudaStream_t st) {
  if (stride == nchan) {
    CIP_CHECK_CUDA(cudaMemcpyAsync(dst, src, elem * (size_t)nrow * nchan, cudaMemcpyHostToDevice, st));
  } else {
    CIP_CHECK_CUDA(cudaMemcpy2DAsync(dst, elem * nchan, src, elem * stride, elem * nchan, nrow, cudaMemcpyHostToDevice, st));
  }
  return CIP_OK;
}

// Host inputs are copied to context-owned device buffers (rows compacted to stride nchan);
// device inputs are used in place.  See common.h for the streaming contract.
cip_status stage_inputs(cip_ctx* ctx, const cip_invert_args* a, DevInputs* d, VisFeed* feed, bool reuse_small,
                        double* ms_h2d) {
  const int64_t nrow = a->nrow, nchan = a->nchan;
  const int64_t vs = a->vis_row_stride ? a->vis_row_stride : nchan;
  const int64_t ws = a->wgt_row_stride ? a->wgt_row_stride : nchan;
  d->nrow = nrow;
  d->nchan = nchan;
  d->w_lo = a->w_hi > 0 ? a->w_lo : 0.0;
  d->w_hi = a->w_hi > 0 ? a->w_hi : 0.0;
  d->wide = a->vis_c128 ? 1 : 0;
  const size_t vis_elem = d->wide ? sizeof(double2) : sizeof(float2), wgt_elem = d->wide ? sizeof(double) : sizeof(float);
  if (ms_h2d) *ms_h2d = 0.0;
  const bool reuse = reuse_small && ctx->staged_valid && a->memspace == 0 && ctx->staged_key[0] == a->uvw &&
                     ctx->staged_key[1] == a->freq && ctx->staged_key[2] == a->wgt && ctx->staged_key[3] == a->mask &&
                     ctx->staged_dims[0] == nrow && ctx->staged_dims[1] == nchan && ctx->staged_dims[2] == ws && ctx->staged_dims[3] == d->wide;
  ctx->staged_valid = false;
  if (a->memspace == 1) {
    d->uvw = a->uvw;
    d->freq = a->freq;
    d->vis = static_cast<const float2*>(a->vis);
    d->wgt = static_cast<const float*>(a->wgt);
    d->mask = a->mask;
    d->vstride = vs;
    d->wstride = ws;
    d->voff = a->row_vis_offset;
    d->ch0 = a->row_chan_start;
    d->ch1 = a->row_chan_stop;
    d->nvis = a->row_vis_offset ? a->nvis : 0;
    if (feed) {
      // Device inputs are all there; they are still cut into row chunks so that the binning of chunk c+1
      // (context stream, bound by memory latency) overlaps the gridding of chunk c (grid stream, bound by
      // the FMA and shared-memory pipes).  More chunks hide more of the binning but multiply the work
      // items (one sub-grid flush each).
      const size_t vis_bytes = (size_t)nrow * nchan * vis_elem;
      int nchunks = 1;  // measured on B200 (configs[1]): 1 chunk 22.8 ms, 2 chunks 23.5 ms, 3 chunks 24.2 ms -- next
                        // to the gridder's CTAs the binning kernels get half their occupancy, so the default stays 1
      (void)vis_bytes;
      if (const char* e = getenv("CIP_DEVICE_CHUNKS")) nchunks = std::min(kMaxChunks, std::max(1, atoi(e)));
      nchunks = (int)std::min<int64_t>(nchunks, std::max<int64_t>(nrow, 1));
      if (a->row_vis_offset) nchunks = 1;
      feed->nchunks = nchunks;
      for (int c = 0; c <= nchunks; ++c) feed->row0[c] = nrow * c / nchunks;
      for (int c = 0; c < nchunks; ++c) feed->ready[c] = nullptr;  // nothing to wait for
    }
    return CIP_OK;
  }
  cudaStream_t st = ctx->stream;
  ctx->t_stage_begin = now_ms();
  if (a->row_vis_offset) {
    // jagged rows from host memory: everything in one piece (no chunked stream)
    const size_t nvis = (size_t)a->nvis;
    CIP_TRY(ctx->in_uvw.ensure(sizeof(double) * 3 * (size_t)nrow + 8));
    CIP_TRY(ctx->in_freq.ensure(sizeof(double) * (size_t)nchan + 8));
    CIP_TRY(ctx->in_rows.ensure(16 * (size_t)nrow + 64));
    if (a->wgt) CIP_TRY(ctx->in_wgt.ensure(wgt_elem * nvis + 8));
    if (a->mask) CIP_TRY(ctx->in_mask.ensure(nvis + 8));
    if (feed) CIP_TRY(ctx->in_vis.ensure(vis_elem * nvis + 8));
    int64_t* d_off = ctx->in_rows.as<int64_t>();
    int32_t* d_c0 = reinterpret_cast<int32_t*>(d_off + nrow);
    int32_t* d_c1 = d_c0 + nrow;
    d->uvw = ctx->in_uvw.as<double>();
    d->freq = ctx->in_freq.as<double>();
    d->vis = feed ? ctx->in_vis.as<float2>() : nullptr;
    d->wgt = a->wgt ? ctx->in_wgt.as<float>() : nullptr;
    d->mask = a->mask ? ctx->in_mask.as<uint8_t>() : nullptr;
    d->vstride = d->wstride = nchan;
    d->voff = d_off;
    d->ch0 = d_c0;
    d->ch1 = d_c1;
    d->nvis = a->nvis;
    if (nrow) {
      CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_uvw.p, a->uvw, sizeof(double) * 3 * nrow, cudaMemcpyHostToDevice, st));
      CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_freq.p, a->freq, sizeof(double) * nchan, cudaMemcpyHostToDevice, st));
      CIP_CHECK_CUDA(cudaMemcpyAsync(d_off, a->row_vis_offset, sizeof(int64_t) * nrow, cudaMemcpyHostToDevice, st));
      CIP_CHECK_CUDA(cudaMemcpyAsync(d_c0, a->row_chan_start, sizeof(int32_t) * nrow, cudaMemcpyHostToDevice, st));
      CIP_CHECK_CUDA(cudaMemcpyAsync(d_c1, a->row_chan_stop, sizeof(int32_t) * nrow, cudaMemcpyHostToDevice, st));
      if (a->wgt && nvis) CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_wgt.p, a->wgt, wgt_elem * nvis, cudaMemcpyHostToDevice, st));
      if (a->mask && nvis) CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_mask.p, a->mask, nvis, cudaMemcpyHostToDevice, st));
      if (feed && nvis) CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_vis.p, a->vis, vis_elem * nvis, cudaMemcpyHostToDevice, st));
    }
    CIP_CHECK_CUDA(cudaEventRecord(ctx->small_ev, st));
    if (feed) {
      feed->nchunks = 1;
      feed->row0[0] = 0;
      feed->row0[1] = nrow;
      feed->ready[0] = nullptr;  // same stream: already ordered
    }
    if (ms_h2d) {
      CIP_CHECK_CUDA(cudaEventSynchronize(ctx->small_ev));
      *ms_h2d = now_ms() - ctx->t_stage_begin;
    }
    return CIP_OK;
  }
  CIP_TRY(ctx->in_uvw.ensure(sizeof(double) * 3 * (size_t)nrow + 8));
  CIP_TRY(ctx->in_freq.ensure(sizeof(double) * (size_t)nchan + 8));
  if (a->wgt) CIP_TRY(ctx->in_wgt.ensure(wgt_elem * (size_t)nrow * nchan + 8));
  if (a->mask) CIP_TRY(ctx->in_mask.ensure((size_t)nrow * nchan + 8));
  if (feed) CIP_TRY(ctx->in_vis.ensure(vis_elem * (size_t)nrow * nchan + 8));
  d->uvw = ctx->in_uvw.as<double>();
  d->freq = ctx->in_freq.as<double>();
  d->vis = feed ? ctx->in_vis.as<float2>() : nullptr;
  d->vstride = nchan;
  d->wstride = nchan;
  d->wgt = a->wgt ? ctx->in_wgt.as<float>() : nullptr;
  d->mask = a->mask ? ctx->in_mask.as<uint8_t>() : nullptr;
  if (nrow && nchan && !reuse) {
    CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_uvw.p, a->uvw, sizeof(double) * 3 * nrow, cudaMemcpyHostToDevice, st));
    CIP_CHECK_CUDA(cudaMemcpyAsync(ctx->in_freq.p, a->freq, sizeof(double) * nchan, cudaMemcpyHostToDevice, st));
    if (a->wgt) CIP_TRY(copy_rows(ctx->in_wgt.p, a->wgt, wgt_elem, nrow, nchan, ws, st));
    if (a->mask) CIP_TRY(copy_rows(ctx->in_mask.p, a->mask, 1, nrow, nchan, ws, st));
  }
  CIP_CHECK_CUDA(cudaEventRecord(ctx->small_ev, st));
  if (feed) {
    // visibilities: row chunks of >= 64 MB on the copy stream, behind the small inputs so that the scan
    // and the planning can start as early as possible
    const size_t vis_bytes = vis_elem * (size_t)nrow * nchan;
    int nchunks = (int)std::min<size_t>(kMaxChunks, std::max<size_t>(1, vis_bytes >> 26));
    if (const char* e = getenv("CIP_STREAM_CHUNKS")) nchunks = std::min(kMaxChunks, std::max(1, atoi(e)));  // tests
    nchunks = (int)std::min<int64_t>(nchunks, std::max<int64_t>(nrow, 1));
    feed->nchunks = nchunks;
    CIP_CHECK_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->small_ev, 0));
    // The compute keeps up with the upload, so what the caller waits for after the last byte has arrived is
    // the binning and gridding of the LAST chunk: the final three chunks taper (1/2, 1/4, 1/8 of the others).
    double wsum = 0.0, wacc = 0.0, wgt_c[kMaxChunks];
    for (int c = 0; c < nchunks; ++c) {
      const int from_end = nchunks - 1 - c;
      wgt_c[c] = (nchunks >= 4 && from_end < 3 && !getenv("CIP_STREAM_CHUNKS")) ? 1.0 / (double)(1 << (3 - from_end)) : 1.0;
      wsum += wgt_c[c];
    }
    int64_t r_prev = 0;
    for (int c = 0; c < nchunks; ++c) {
      wacc += wgt_c[c];
      const int64_t r0 = r_prev, r1 = c == nchunks - 1 ? nrow : std::max(r_prev, (int64_t)((double)nrow * (wacc / wsum)));
      r_prev = r1;
      feed->row0[c] = r0;
      feed->row0[c + 1] = r1;
      if (r1 > r0 && nchan)
        CIP_TRY(copy_rows(static_cast<char*>(ctx->in_vis.p) + vis_elem * r0 * nchan, static_cast<const char*>(a->vis) + vis_elem * r0 * vs,
                          vis_elem, r1 - r0, nchan, vs, ctx->copy_stream));
      CIP_CHECK_CUDA(cudaEventRecord(ctx->chunk_ev[c], ctx->copy_stream));
      feed->ready[c] = ctx->chunk_ev[c];
    }
  }
  if (ms_h2d) {  // host clock: see StageTimer in invert.cu for why no timing event is recorded here
    CIP_CHECK_CUDA(cudaEventSynchronize(ctx->small_ev));
    *ms_h2d = now_ms() - ctx->t_stage_begin;
  }
  ctx->staged_key[0] = a->uvw;
  ctx->staged_key[1] = a->freq;
  ctx->staged_key[2] = a->wgt;
  ctx->staged_key[3] = a->mask;
  ctx->staged_dims[0] = nrow;
  ctx->staged_dims[1] = nchan;
  ctx->staged_dims[2] = ws;
  ctx->staged_dims[3] = d->wide;
  return CIP_OK;
}

__global__ void small_copy_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, int nwords) {
  for (int i = threadIdx.x; i < nwords; i += blockDim.x) dst[i] = src[i];
  __threadfence_system();
}

cip_status fetch_small(cip_ctx* ctx, const void* dev, size_t bytes, void* host_out) {
  CIP_REQUIRE(bytes % 4 == 0 && bytes <= kSmallHostBytes, "fetch_small: bad size %zu", bytes);
  small_copy_kernel<<<1, 256, 0, ctx->stream>>>(static_cast<const uint32_t*>(dev),
                                               static_cast<uint32_t*>(ctx->scan_host_dev), (int)(bytes / 4));
  CIP_CHECK_CUDA(cudaGetLastError());
  CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
  memcpy(host_out, ctx->scan_host, bytes);
  return CIP_OK;
}

cip_status wait_feed(cip_ctx* ctx, const VisFeed& feed) {
  for (int c = 0; c < feed.nchunks; ++c)
    if (feed.ready[c]) CIP_CHECK_CUDA(cudaStreamWaitEvent(ctx->stream, feed.ready[c], 0));
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
  DeviceScope scope(device);
  CIP_CHECK_CUDA(scope.err);
  cip_ctx* ctx = new cip_ctx();
  ctx->device = device;
  cudaDeviceProp prop{};
  CIP_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
  ctx->sm_count = prop.multiProcessorCount;
  ctx->smem_optin = prop.sharedMemPerBlockOptin;
  CIP_CHECK_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  CIP_CHECK_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  {
    int lo = 0, hi = 0;  // the gridder yields to the (short, latency-bound) binning kernels of the next chunk
    CIP_CHECK_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CIP_CHECK_CUDA(cudaStreamCreateWithPriority(&ctx->grid_stream, cudaStreamNonBlocking, getenv("CIP_NO_PRIO") ? 0 : lo));
  }
  for (int s = 0; s < 2; ++s) {
    CIP_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->binned_ev[s], cudaEventDisableTiming));
    CIP_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->gridded_ev[s], cudaEventDisableTiming));
  }
  CIP_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->zero_ev, cudaEventDisableTiming));
  for (auto& row : ctx->chunk_t)
    for (auto& ev : row) CIP_CHECK_CUDA(cudaEventCreate(&ev));
  for (auto& ev : ctx->ev) CIP_CHECK_CUDA(cudaEventCreate(&ev));
  for (auto& ev : ctx->chunk_ev) CIP_CHECK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  CIP_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->sync_ev, cudaEventDisableTiming));
  CIP_CHECK_CUDA(cudaEventCreateWithFlags(&ctx->small_ev, cudaEventDisableTiming));
  CIP_CHECK_CUDA(cudaHostAlloc(&ctx->scan_host, kSmallHostBytes, cudaHostAllocMapped));
  CIP_CHECK_CUDA(cudaHostGetDevicePointer(&ctx->scan_host_dev, ctx->scan_host, 0));
  CIP_CHECK_CUDA(cudaHostAlloc(&ctx->flags_host, 256, cudaHostAllocMapped));
  CIP_CHECK_CUDA(cudaHostGetDevicePointer(&ctx->flags_host_dev, ctx->flags_host, 0));
  memset(ctx->flags_host, 0, 256);
  CIP_TRY(ctx->call_flags.ensure(64));
  CIP_CHECK_CUDA(cudaMemset(ctx->call_flags.p, 0, 64));
  *out = ctx;
  return CIP_OK;
}

extern "C" cip_status cip_ctx_destroy(cip_ctx* ctx) {
  if (!ctx) return CIP_OK;
  DeviceScope scope(ctx->device);
  if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->grid_stream) cudaStreamSynchronize(ctx->grid_stream);
  for (auto& kv : ctx->fft_plans) cufftDestroy(kv.second.handle);
  cip::pfft_release(ctx);
  DevBuf* bufs[] = {&ctx->in_uvw, &ctx->in_freq, &ctx->in_vis, &ctx->in_wgt, &ctx->in_mask, &ctx->in_rows, &ctx->out_img,
                    &ctx->scan_out, &ctx->grid, &ctx->image, &ctx->tables, &ctx->fft_t, &ctx->t_uvw, &ctx->t_freq, &ctx->t_vis,
                    &ctx->call_flags};
  for (auto& w : ctx->work) w.release();
  ctx->twork.release();
  for (DevBuf* b : bufs) b->release();
  if (ctx->scan_host) cudaFreeHost(ctx->scan_host);
  if (ctx->flags_host) cudaFreeHost(ctx->flags_host);
  for (auto& ev : ctx->ev)
    if (ev) cudaEventDestroy(ev);
  for (auto& ev : ctx->chunk_ev)
    if (ev) cudaEventDestroy(ev);
  for (int s = 0; s < 2; ++s) {
    if (ctx->binned_ev[s]) cudaEventDestroy(ctx->binned_ev[s]);
    if (ctx->gridded_ev[s]) cudaEventDestroy(ctx->gridded_ev[s]);
  }
  if (ctx->zero_ev) cudaEventDestroy(ctx->zero_ev);
  for (auto& row : ctx->chunk_t)
    for (auto& ev : row)
      if (ev) cudaEventDestroy(ev);
  if (ctx->grid_stream) cudaStreamDestroy(ctx->grid_stream);
  if (ctx->small_ev) cudaEventDestroy(ctx->small_ev);
  if (ctx->sync_ev) cudaEventDestroy(ctx->sync_ev);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return CIP_OK;
}

extern "C" void* cip_ctx_stream(cip_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

extern "C" int64_t cip_ctx_launches(cip_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" int cip_ctx_device(cip_ctx* ctx) { return ctx ? ctx->device : -1; }
extern "C" cip_status cip_current_device(int* device) {
  CIP_REQUIRE(device != nullptr, "device output is NULL");
  CIP_CHECK_CUDA(cudaGetDevice(device));
  return CIP_OK;
}

extern "C" cip_status cip_ctx_synchronize(cip_ctx* ctx) {
  CIP_REQUIRE(ctx != nullptr, "ctx is NULL");
  CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
  return CIP_OK;
}
