// G1 scan: |w| range over the active samples, active count, and the weight sum that the
// reference takes on the host (effective_weights.sum(), src/ska_sdp_cip/invert.py:184).
// A sample is active when wgt * mask != 0.  (ducc0 also drops samples whose visibility is exactly
// zero; they add nothing to the image either way, so here they merely may widen the planned w range.
// Not reading the visibilities lets the planning start while they are still in flight from the
// host.)  One pass over wgt/mask (HBM-bound: 4-5 B per sample), block reduction, 4 atomics per CTA.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "geom.cuh"

namespace cip {

struct ScanOut {
  unsigned long long wmin_bits, wmax_bits;  // non-negative doubles order like their bit patterns
  unsigned long long nactive;
  double sum_wgt;
};

struct ScanAcc {
  double lo = 1e300, hi = 0.0, sw = 0.0;
  unsigned long long na = 0;
  double w_lo = 0.0, w_hi = 0.0;
  __device__ __forceinline__ void add(double w, bool masked, double aw) {
    if (masked || (w_hi > 0.0 && !(aw >= w_lo && aw < w_hi))) w = 0.0;
    sw += w;
    if (w != 0.0) {
      lo = fmin(lo, aw);
      hi = fmax(hi, aw);
      ++na;
    }
  }
};

__device__ __forceinline__ void scan_finish(ScanAcc a, ScanOut* out) {
  double lo = a.lo, hi = a.hi, sw = a.sw;
  unsigned long long na = a.na;
  for (int o = 16; o; o >>= 1) {
    lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    sw += __shfl_xor_sync(0xffffffffu, sw, o);
    na += __shfl_xor_sync(0xffffffffu, na, o);
  }
  __shared__ double s_lo[8], s_hi[8], s_sw[8];
  __shared__ unsigned long long s_na[8];
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_lo[warp] = lo; s_hi[warp] = hi; s_sw[warp] = sw; s_na[warp] = na; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < 8; ++i) { lo = fmin(lo, s_lo[i]); hi = fmax(hi, s_hi[i]); sw += s_sw[i]; na += s_na[i]; }
    if (na) {
      atomicMin(&out->wmin_bits, (unsigned long long)__double_as_longlong(lo));
      atomicMax(&out->wmax_bits, (unsigned long long)__double_as_longlong(hi));
      atomicAdd(&out->nactive, na);
    }
    atomicAdd(&out->sum_wgt, sw);
  }
}

// general layout: one sample per thread and iteration
__global__ void __launch_bounds__(256) scan_kernel(DevInputs in, int do_wstacking, ScanOut* out) {
  const int64_t nchan = in.nchan;
  const int64_t total = in.nrow * nchan;
  ScanAcc acc;
  acc.w_lo = in.w_lo;
  acc.w_hi = in.w_hi;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = k / nchan, c = k - r * nchan;
    double aw = do_wstacking ? fabs(in.uvw[3 * r + 2]) * (in.freq[c] / kSpeedOfLight) : 0.0;
    acc.add(sample_weight(in, r * in.wstride + c), false, aw);
  }
  scan_finish(acc, out);
}

// jagged rows (and any other layout): one warp per row, lanes stride the row's channels
__global__ void __launch_bounds__(256) scan_rows_kernel(DevInputs in, int do_wstacking, ScanOut* out) {
  ScanAcc acc;
  acc.w_lo = in.w_lo;
  acc.w_hi = in.w_hi;
  const int lane = threadIdx.x & 31;
  for (int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < in.nrow; row += ((int64_t)gridDim.x * blockDim.x) >> 5) {
    const RowSpan sp = row_span(in, row);
    const double wm = do_wstacking ? fabs(in.uvw[3 * row + 2]) : 0.0;
    for (int c = sp.c_begin + lane; c < sp.c_end; c += 32) {
      acc.add(sample_weight(in, sp.wbase + c), false, wm * (in.freq[c] / kSpeedOfLight));
    }
  }
  scan_finish(acc, out);
}

// nchan, strides multiples of 4 and 16-byte aligned pointers: four channels per thread and
// iteration with 128-bit loads (2 x float4 vis, float4 wgt, one 32-bit word of mask)
// (frequency / c per channel is tabulated in shared memory: a float64 division per sample made this kernel
// compute-bound at a third of the HBM rate)
template <typename Index>
__global__ void __launch_bounds__(256) scan_kernel_v4(DevInputs in, int do_wstacking, ScanOut* out) {
  extern __shared__ double s_finv[];
  for (int c = threadIdx.x; c < in.nchan; c += blockDim.x) s_finv[c] = in.freq[c] / kSpeedOfLight;
  __syncthreads();
  const Index nq = (Index)(in.nchan >> 2);
  const Index total = (Index)(in.nrow * (int64_t)nq);
  ScanAcc acc;
  acc.w_lo = in.w_lo;
  acc.w_hi = in.w_hi;
  for (Index k = (Index)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (Index)gridDim.x * blockDim.x) {
    const Index r = k / nq;
    const int c = (int)(k - r * nq) << 2;
    float4 w = make_float4(1.f, 1.f, 1.f, 1.f);
    if (in.wgt) w = __ldg(reinterpret_cast<const float4*>(in.wgt + (int64_t)r * in.wstride + c));
    uint32_t mk = 0x01010101u;
    if (in.mask) mk = __ldg(reinterpret_cast<const uint32_t*>(in.mask + (int64_t)r * in.wstride + c));
    double aw0 = 0.0, aw1 = 0.0, aw2 = 0.0, aw3 = 0.0;
    if (do_wstacking) {
      const double wm = fabs(in.uvw[3 * (int64_t)r + 2]);
      aw0 = wm * s_finv[c];
      aw1 = wm * s_finv[c + 1];
      aw2 = wm * s_finv[c + 2];
      aw3 = wm * s_finv[c + 3];
    }
    acc.add(w.x, (mk & 0xffu) == 0, aw0);
    acc.add(w.y, (mk & 0xff00u) == 0, aw1);
    acc.add(w.z, (mk & 0xff0000u) == 0, aw2);
    acc.add(w.w, (mk & 0xff000000u) == 0, aw3);
  }
  scan_finish(acc, out);
}

__global__ void scan_init_kernel(ScanOut* out) {
  out->wmin_bits = 0x7ff0000000000000ull;  // +inf
  out->wmax_bits = 0;
  out->nactive = 0;
  out->sum_wgt = 0.0;
}

static double bits_to_double(unsigned long long b) {
  double d;
  memcpy(&d, &b, sizeof(d));
  return d;
}

// |w| histogram of the active samples (the slab filter of the inputs, if any, applies too).
constexpr int kMaxWBins = 4096;
__global__ void __launch_bounds__(256) whist_kernel(DevInputs in, double wmin, double scale, int nbins, unsigned long long* __restrict__ counts) {
  __shared__ uint32_t s_cnt[kMaxWBins];
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) s_cnt[b] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  for (int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < in.nrow; row += ((int64_t)gridDim.x * blockDim.x) >> 5) {
    const RowSpan sp = row_span(in, row);
    const double wm = fabs(in.uvw[3 * row + 2]);
    for (int c = sp.c_begin + lane; c < sp.c_end; c += 32) {
      if (sample_weight(in, sp.wbase + c) == 0.0) continue;
      const double aw = wm * (in.freq[c] / kSpeedOfLight);
      if (in.w_hi > 0.0 && !(aw >= in.w_lo && aw < in.w_hi)) continue;
      int b = (int)((aw - wmin) * scale);
      b = b < 0 ? 0 : (b >= nbins ? nbins - 1 : b);
      atomicAdd(&s_cnt[b], 1u);
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nbins; b += blockDim.x)
    if (s_cnt[b]) atomicAdd(&counts[b], (unsigned long long)s_cnt[b]);
}

cip_status run_w_histogram(cip_ctx* ctx, const DevInputs& in, double wmin, double wmax, int nbins, int64_t* counts_out) {
  CIP_REQUIRE(nbins >= 1 && nbins <= kMaxWBins, "nbins must be in [1, %d]", kMaxWBins);
  CIP_REQUIRE(wmax >= wmin, "bad |w| range");
  CIP_TRY(ctx->scan_out.ensure(sizeof(unsigned long long) * kMaxWBins));
  cudaStream_t st = ctx->stream;
  unsigned long long* d = ctx->scan_out.as<unsigned long long>();
  CIP_CHECK_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long) * nbins, st));
  const int64_t total = in.nrow * in.nchan;
  if (total > 0) {
    const double scale = wmax > wmin ? nbins / (wmax - wmin) : 0.0;
    const int blocks = (int)std::min<int64_t>((in.nrow * 32 + 255) / 256, (int64_t)ctx->sm_count * 8);
    whist_kernel<<<blocks, 256, 0, st>>>(in, wmin, scale, nbins, d);
    ctx->launches++;
    CIP_CHECK_CUDA(cudaGetLastError());
  }
  std::vector<unsigned long long> h(nbins);
  CIP_CHECK_CUDA(cudaMemcpyAsync(h.data(), d, sizeof(unsigned long long) * nbins, cudaMemcpyDeviceToHost, st));
  CIP_CHECK_CUDA(cudaStreamSynchronize(st));
  for (int b = 0; b < nbins; ++b) counts_out[b] += (int64_t)h[b];
  return CIP_OK;
}

// ---- the |w| histogram without any allocation (cip_w_histogram_static) ----------------------------------------
// Scratch of the histogram kernel: the per-bin counters, here in a file-scope array instead of a context buffer.  Every
// GPU that loads the library holds these 32 KB until exit; all calls on a device share them.
static_assert(kMaxWBins == CIP_W_HISTOGRAM_STATIC_MAX_BINS, "cip.h states the limit");
static __device__ __align__(256) unsigned long long g_whist_scratch[kMaxWBins];

// Device address of the array on the current device.  In host code `g_whist_scratch` is only a stand-in for the
// symbol, not a device address: it is resolved once per device, with that device current.
static cip_status whist_scratch_ptr(int device, unsigned long long** out) {
  constexpr int kMaxDevices = 64;
  static unsigned long long* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  CIP_REQUIRE(device >= 0 && device < kMaxDevices, "device index %d out of range", device);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[device]) {
    unsigned long long* p = nullptr;
    CIP_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_whist_scratch));
    ptrs[device] = p;
  }
  *out = ptrs[device];
  return CIP_OK;
}

cip_status run_w_histogram_static(cip_ctx* ctx, const DevInputs& in, double wmin, double wmax, int nbins, int64_t* counts_out) {
  CIP_REQUIRE(nbins >= 1, "nbins must be positive");
  CIP_REQUIRE(nbins <= kMaxWBins, "nbins = %d exceeds the %d bins of the static scratch array of cip_w_histogram_static", nbins,
              kMaxWBins);
  CIP_REQUIRE(wmax >= wmin, "bad |w| range");
  unsigned long long* d = nullptr;
  CIP_TRY(whist_scratch_ptr(ctx->device, &d));
  cudaStream_t st = ctx->stream;
  // the array keeps the counts of the previous call: clear the bins this call uses
  CIP_CHECK_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long) * nbins, st));
  if (in.nrow * in.nchan > 0) {
    const double scale = wmax > wmin ? nbins / (wmax - wmin) : 0.0;
    const int blocks = (int)std::min<int64_t>((in.nrow * 32 + 255) / 256, (int64_t)ctx->sm_count * 8);
    whist_kernel<<<blocks, 256, 0, st>>>(in, wmin, scale, nbins, d);
    ctx->launches++;
    CIP_CHECK_CUDA(cudaGetLastError());
  }
  unsigned long long h[kMaxWBins];  // 32 KB of stack: no host allocation either
  CIP_CHECK_CUDA(cudaMemcpyAsync(h, d, sizeof(unsigned long long) * nbins, cudaMemcpyDeviceToHost, st));
  CIP_CHECK_CUDA(cudaStreamSynchronize(st));
  for (int b = 0; b < nbins; ++b) counts_out[b] += (int64_t)h[b];
  return CIP_OK;
}

cip_status run_scan(cip_ctx* ctx, const DevInputs& in, int do_wstacking, cip_scan_result* out) {
  NvtxRange nvtx("cip:scan");
  CIP_TRY(ctx->scan_out.ensure(sizeof(ScanOut)));
  cudaStream_t st = ctx->stream;
  scan_init_kernel<<<1, 1, 0, st>>>(ctx->scan_out.as<ScanOut>());
  int64_t total = in.voff ? in.nvis : in.nrow * in.nchan;
  if (total > 0 && in.voff) {
    const int blocks = (int)std::min<int64_t>((in.nrow * 32 + 255) / 256, (int64_t)ctx->sm_count * 8);
    scan_rows_kernel<<<blocks, 256, 0, st>>>(in, do_wstacking, ctx->scan_out.as<ScanOut>());
    ctx->launches++;
    CIP_CHECK_CUDA(cudaGetLastError());
  } else if (total > 0) {
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool v4 = !in.wide && in.nchan % 4 == 0 && in.wstride % 4 == 0 && al16(in.wgt) && al16(in.mask) && in.nchan <= 4096;
    if (v4) {
      int blocks = (int)std::min<int64_t>((total / 4 + 255) / 256, (int64_t)ctx->sm_count * 8);
      const size_t smem = sizeof(double) * in.nchan;
      if (total / 4 < ((int64_t)1 << 31))  // 32-bit index arithmetic (the row index costs a division per thread)
        scan_kernel_v4<uint32_t><<<blocks, 256, smem, st>>>(in, do_wstacking, ctx->scan_out.as<ScanOut>());
      else
        scan_kernel_v4<int64_t><<<blocks, 256, smem, st>>>(in, do_wstacking, ctx->scan_out.as<ScanOut>());
    } else {
      int blocks = (int)std::min<int64_t>((total + 255) / 256, (int64_t)ctx->sm_count * 8);
      scan_kernel<<<blocks, 256, 0, st>>>(in, do_wstacking, ctx->scan_out.as<ScanOut>());
    }
    ctx->launches++;
    CIP_CHECK_CUDA(cudaGetLastError());
  }
  ScanOut res;
  CIP_TRY(fetch_small(ctx, ctx->scan_out.p, sizeof(ScanOut), &res));
  const ScanOut* host = &res;
  out->nactive = (int64_t)host->nactive;
  out->sum_wgt = host->sum_wgt;
  if (host->nactive) {
    out->wmin = bits_to_double(host->wmin_bits);
    out->wmax = bits_to_double(host->wmax_bits);
  } else {
    out->wmin = out->wmax = 0.0;
  }
  return CIP_OK;
}

}  // namespace cip

// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (see include/cip.h).
extern "C" cip_status cip_w_histogram_static(cip_ctx* ctx, const cip_invert_args* args, double wmin, double wmax, int nbins,
                                             int64_t* counts_out) {
  using namespace cip;
  CIP_REQUIRE(ctx && counts_out, "NULL argument");
  CIP_TRY(validate_args(args));
  CIP_REQUIRE(args->memspace == 1, "cip_w_histogram_static takes device inputs only (memspace 1): staging host arrays allocates");
  CIP_ENTER_DEVICE(ctx);
  DevInputs in;
  CIP_TRY(stage_inputs(ctx, args, &in, nullptr, false, nullptr));  // device inputs: pointers are taken as they are
  return run_w_histogram_static(ctx, in, wmin, wmax, nbins, counts_out);
}
