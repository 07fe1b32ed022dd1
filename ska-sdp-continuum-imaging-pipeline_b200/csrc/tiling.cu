// Kernel 1 (uvw_tiling side): bit-exact UVW tile binning, T1-T5 of SURVEY.md section 8a.
//
// Replaces create_uvw_tile_mapping / create_uvw_tile_mapping_sequential /
// _find_all_constant_tile_index_subarrays / merge_tile_mappings
// (src/ska_sdp_cip/uvw_tiling/tiling_plan.py:29-181) and the gather of
// Tile._from_jagged_visibilities_slice (src/ska_sdp_cip/uvw_tiling/tile.py:83-115).
//
//   idx[ch, k] = (int64) floor( (freq[ch] / c) * (uvw[row, k] / tile[k]) + 0.5 )
// evaluated with the same three separately rounded float64 operations numpy performs
// (__ddiv_rn, __dmul_rn, __dadd_rn: no FMA contraction), run-length encoded along the channel
// axis per row (one warp per row, ballot), grouped by tile with a stable radix sort so slices
// stay in (irow, chan_start) order, and the tiles put in first-appearance order -- the order of
// the reference's dict.
#include <algorithm>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cmath>

#include "common.h"

struct cip_tile_result {
  int64_t n_tiles = 0, n_slices = 0, nrow = 0, nchan = 0, row_offset = 0;
  cip::DevBuf tile_coords;   // (n_tiles,3) int64, first-appearance order
  cip::DevBuf tile_offsets;  // (n_tiles+1,) int64
  cip::DevBuf slice_irow;    // (n_slices,) int64 (includes row_offset)
  cip::DevBuf slice_ch0, slice_ch1;  // (n_slices,) int32
};

namespace cip {

namespace {

struct Idx3 { long long u, v, w; };

__device__ __forceinline__ Idx3 tile_index(double um, double vm, double wm, double freq, double tu, double tv, double tw) {
  const double winv = __ddiv_rn(freq, kSpeedOfLight);
  Idx3 r;
  r.u = (long long)floor(__dadd_rn(__dmul_rn(winv, __ddiv_rn(um, tu)), 0.5));
  r.v = (long long)floor(__dadd_rn(__dmul_rn(winv, __ddiv_rn(vm, tv)), 0.5));
  r.w = (long long)floor(__dadd_rn(__dmul_rn(winv, __ddiv_rn(wm, tw)), 0.5));
  return r;
}

struct Range { long long lo[3], hi[3]; };

template <int PASS>
__global__ void __launch_bounds__(256) tile_runs_kernel(const double* __restrict__ uvw, const double* __restrict__ freq,
                                                        int64_t nrow, int64_t nchan, double tu, double tv, double tw,
                                                        int64_t* __restrict__ row_n, const int64_t* __restrict__ row_off,
                                                        int64_t* __restrict__ s_row, int32_t* __restrict__ s_ch0,
                                                        long long* __restrict__ s_idx /* (n,3) */, Range* range) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= nrow) return;
  const double um = uvw[3 * row], vm = uvw[3 * row + 1], wm = uvw[3 * row + 2];
  Idx3 carry{0, 0, 0};
  bool have_carry = false;
  int64_t n = 0;
  const int64_t base = PASS ? row_off[row] : 0;
  long long lo[3] = {LLONG_MAX, LLONG_MAX, LLONG_MAX}, hi[3] = {LLONG_MIN, LLONG_MIN, LLONG_MIN};
  for (int64_t c0 = 0; c0 < nchan; c0 += 32) {
    const int64_t c = c0 + lane;
    Idx3 id{0, 0, 0};
    if (c < nchan) id = tile_index(um, vm, wm, freq[c], tu, tv, tw);
    Idx3 pv;
    pv.u = __shfl_up_sync(0xffffffffu, id.u, 1);
    pv.v = __shfl_up_sync(0xffffffffu, id.v, 1);
    pv.w = __shfl_up_sync(0xffffffffu, id.w, 1);
    bool differs = id.u != pv.u || id.v != pv.v || id.w != pv.w;
    if (lane == 0) differs = !have_carry || id.u != carry.u || id.v != carry.v || id.w != carry.w;
    const bool start = c < nchan && differs;
    const unsigned m = __ballot_sync(0xffffffffu, start);
    if (PASS && start) {
      int64_t pos = base + n + __popc(m & ((1u << lane) - 1));
      s_row[pos] = row;
      s_ch0[pos] = (int32_t)c;
      s_idx[3 * pos] = id.u;
      s_idx[3 * pos + 1] = id.v;
      s_idx[3 * pos + 2] = id.w;
      lo[0] = min(lo[0], id.u); hi[0] = max(hi[0], id.u);
      lo[1] = min(lo[1], id.v); hi[1] = max(hi[1], id.v);
      lo[2] = min(lo[2], id.w); hi[2] = max(hi[2], id.w);
    }
    n += __popc(m);
    carry.u = __shfl_sync(0xffffffffu, id.u, 31);
    carry.v = __shfl_sync(0xffffffffu, id.v, 31);
    carry.w = __shfl_sync(0xffffffffu, id.w, 31);
    have_carry = true;
  }
  if (!PASS && lane == 0) row_n[row] = n;
  if (PASS) {
    for (int k = 0; k < 3; ++k) {
      for (int o = 16; o; o >>= 1) {
        lo[k] = min(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
        hi[k] = max(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
      }
      if (lane == 0 && lo[k] <= hi[k]) {
        atomicMin(&range->lo[k], lo[k]);
        atomicMax(&range->hi[k], hi[k]);
      }
    }
  }
}

__global__ void slice_finish_kernel(const int64_t* __restrict__ s_row, const int32_t* __restrict__ s_ch0,
                                    const long long* __restrict__ s_idx, int64_t n, int32_t nchan, Range range,
                                    unsigned long long rv, unsigned long long rw, int32_t* __restrict__ s_ch1,
                                    uint64_t* __restrict__ keys, uint32_t* __restrict__ ids) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t stop = nchan;
  if (i + 1 < n && s_row[i + 1] == s_row[i]) stop = s_ch0[i + 1];
  s_ch1[i] = stop;
  unsigned long long a = (unsigned long long)(s_idx[3 * i] - range.lo[0]);
  unsigned long long b = (unsigned long long)(s_idx[3 * i + 1] - range.lo[1]);
  unsigned long long c = (unsigned long long)(s_idx[3 * i + 2] - range.lo[2]);
  keys[i] = (a * rv + b) * rw + c;
  ids[i] = (uint32_t)i;
}

__global__ void tile_flags_kernel(const uint64_t* __restrict__ keys_sorted, int64_t n, int64_t* __restrict__ flag) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = (i == 0 || keys_sorted[i] != keys_sorted[i - 1]) ? 1 : 0;
}

// per tile (key order): position of its first slice in sorted order and the id of that slice
__global__ void tile_first_kernel(const int64_t* __restrict__ flag, const int64_t* __restrict__ tile_of /* inclusive */,
                                  const uint32_t* __restrict__ ids_sorted, int64_t n, int64_t* __restrict__ first_pos,
                                  uint32_t* __restrict__ first_id, uint32_t* __restrict__ tile_id) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !flag[i]) return;
  int64_t t = tile_of[i] - 1;
  first_pos[t] = i;
  first_id[t] = ids_sorted[i];  // stable sort: the smallest slice id of the tile = its first appearance
  tile_id[t] = (uint32_t)t;
}

// tiles in first-appearance order: sizes and coordinates
__global__ void tile_order_kernel(const uint32_t* __restrict__ tile_sorted /* key-order tile id per output rank */,
                                  const int64_t* __restrict__ first_pos, const uint32_t* __restrict__ ids_sorted,
                                  const long long* __restrict__ s_idx, int64_t n_tiles, int64_t n_slices,
                                  int64_t* __restrict__ sizes, int64_t* __restrict__ coords) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > n_tiles) return;
  if (r == n_tiles) { sizes[r] = 0; return; }
  uint32_t t = tile_sorted[r];
  int64_t b = first_pos[t], e = (int64_t)t + 1 < n_tiles ? first_pos[t + 1] : n_slices;
  sizes[r] = e - b;
  uint32_t id = ids_sorted[b];
  coords[3 * r] = s_idx[3 * (int64_t)id];
  coords[3 * r + 1] = s_idx[3 * (int64_t)id + 1];
  coords[3 * r + 2] = s_idx[3 * (int64_t)id + 2];
}

// one thread block per output tile: copy its slices (already in ascending (irow, ch) order)
__global__ void tile_emit_kernel(const uint32_t* __restrict__ tile_sorted, const int64_t* __restrict__ first_pos,
                                 const int64_t* __restrict__ out_off, const uint32_t* __restrict__ ids_sorted,
                                 const int64_t* __restrict__ s_row, const int32_t* __restrict__ s_ch0,
                                 const int32_t* __restrict__ s_ch1, int64_t row_offset, int64_t* __restrict__ o_row,
                                 int32_t* __restrict__ o_ch0, int32_t* __restrict__ o_ch1) {
  const int64_t r = blockIdx.x;
  const uint32_t t = tile_sorted[r];
  const int64_t b = first_pos[t], o = out_off[r], cnt = out_off[r + 1] - o;
  for (int64_t k = threadIdx.x; k < cnt; k += blockDim.x) {
    uint32_t id = ids_sorted[b + k];
    o_row[o + k] = s_row[id] + row_offset;
    o_ch0[o + k] = s_ch0[id];
    o_ch1[o + k] = s_ch1[id];
  }
}

__global__ void slice_len_kernel(const int32_t* __restrict__ ch0, const int32_t* __restrict__ ch1, int64_t n,
                                 int64_t* __restrict__ len) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= n) len[i] = i < n ? (int64_t)(ch1[i] - ch0[i]) : 0;
}

// one warp per slice
__global__ void __launch_bounds__(256) gather_kernel(const float2* __restrict__ vis, const double* __restrict__ uvw,
                                                     int64_t nchan, int64_t row_offset, const int64_t* __restrict__ o_row,
                                                     const int32_t* __restrict__ o_ch0, const int64_t* __restrict__ off,
                                                     int64_t n, float2* __restrict__ vis_out, double* __restrict__ uvw_out) {
  const int lane = threadIdx.x & 31;
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i >= n) return;
  const int64_t row = o_row[i] - row_offset;
  const int64_t b = off[i], len = off[i + 1] - b;
  const float2* src = vis + row * nchan + o_ch0[i];
  for (int64_t k = lane; k < len; k += 32) vis_out[b + k] = src[k];
  if (lane < 3) uvw_out[3 * i + lane] = uvw[3 * row + lane];
}

inline int nblk(int64_t n, int per) { return (int)((n + per - 1) / per); }

int bits_for(unsigned long long n) {  // bits needed to represent values 0..n-1
  int b = 0;
  while (b < 64 && (n - 1) >> b) ++b;
  return std::max(b, 1);
}

}  // namespace
}  // namespace cip

using namespace cip;

extern "C" cip_status cip_tile_free(cip_ctx* ctx, cip_tile_result* r) {
  (void)ctx;
  if (!r) return CIP_OK;
  r->tile_coords.release();
  r->tile_offsets.release();
  r->slice_irow.release();
  r->slice_ch0.release();
  r->slice_ch1.release();
  delete r;
  return CIP_OK;
}

extern "C" cip_status cip_tile_bin(cip_ctx* ctx, const double* uvw, int64_t nrow, const double* freq, int64_t nchan,
                                   const double tile_size[3], int64_t row_offset, int memspace,
                                   cip_tile_result** result, int64_t* n_tiles, int64_t* n_slices) {
  CIP_REQUIRE(ctx && result && tile_size, "NULL argument");
  CIP_REQUIRE(nrow >= 0 && nchan >= 0 && nrow < ((int64_t)1 << 31) && nchan < ((int64_t)1 << 24), "bad nrow/nchan");
  CIP_REQUIRE(nrow == 0 || (uvw != nullptr), "uvw is NULL");
  CIP_REQUIRE(nchan > 0 || nrow == 0, "channel_freqs is empty (the reference raises IndexError here)");
  CIP_REQUIRE(nchan == 0 || freq != nullptr, "freq is NULL");
  CIP_REQUIRE(memspace == 0 || memspace == 1, "memspace must be 0 or 1");
  for (int k = 0; k < 3; ++k)
    CIP_REQUIRE(tile_size[k] > 0 && tile_size[k] < HUGE_VAL, "tile_size[%d] must be positive and finite", k);
  CIP_ENTER_DEVICE(ctx);
  *result = nullptr;
  cudaStream_t st = ctx->stream;
  cip_tile_result* R = new cip_tile_result();
  R->nrow = nrow;
  R->nchan = nchan;
  R->row_offset = row_offset;
  auto fail = [&](cip_status s) { cip_tile_free(ctx, R); return s; };
#define TTRY(expr) do { cip_status _s = (expr); if (_s != CIP_OK) return fail(_s); } while (0)
#define TCUDA(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { set_error("%s failed: %s", #expr, cudaGetErrorString(_e)); return fail(CIP_ECUDA); } } while (0)
  if (n_tiles) *n_tiles = 0;
  if (n_slices) *n_slices = 0;
  TTRY(R->tile_offsets.ensure(sizeof(int64_t)));
  TCUDA(cudaMemsetAsync(R->tile_offsets.p, 0, sizeof(int64_t), st));
  if (nrow == 0) {
    TCUDA(cudaStreamSynchronize(st));
    *result = R;
    return CIP_OK;
  }
  const double* d_uvw = uvw;
  const double* d_freq = freq;
  if (memspace == 0) {
    TTRY(ctx->t_uvw.ensure(sizeof(double) * 3 * nrow));
    TTRY(ctx->t_freq.ensure(sizeof(double) * nchan));
    TCUDA(cudaMemcpyAsync(ctx->t_uvw.p, uvw, sizeof(double) * 3 * nrow, cudaMemcpyHostToDevice, st));
    TCUDA(cudaMemcpyAsync(ctx->t_freq.p, freq, sizeof(double) * nchan, cudaMemcpyHostToDevice, st));
    d_uvw = ctx->t_uvw.as<double>();
    d_freq = ctx->t_freq.as<double>();
  }
  // --- runs per row ---------------------------------------------------------------------
  TTRY(ctx->twork.row_nruns.ensure(sizeof(int64_t) * (nrow + 1)));
  TTRY(ctx->twork.row_off.ensure(sizeof(int64_t) * (nrow + 1)));
  int64_t* row_n = ctx->twork.row_nruns.as<int64_t>();
  int64_t* row_off = ctx->twork.row_off.as<int64_t>();
  TCUDA(cudaMemsetAsync(row_n + nrow, 0, sizeof(int64_t), st));
  const int rblocks = nblk(nrow * 32, 256);
  tile_runs_kernel<0><<<rblocks, 256, 0, st>>>(d_uvw, d_freq, nrow, nchan, tile_size[0], tile_size[1], tile_size[2],
                                               row_n, nullptr, nullptr, nullptr, nullptr, nullptr);
  size_t tmp = 0;
  TCUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, row_n, row_off, nrow + 1, st));
  TTRY(ctx->twork.cub_tmp.ensure(tmp));
  tmp = ctx->twork.cub_tmp.cap;
  TCUDA(cub::DeviceScan::ExclusiveSum(ctx->twork.cub_tmp.p, tmp, row_n, row_off, nrow + 1, st));
  int64_t ns = 0;
  TCUDA(cudaMemcpyAsync(ctx->scan_host, row_off + nrow, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  TCUDA(cudaStreamSynchronize(st));
  ns = *static_cast<int64_t*>(ctx->scan_host);
  if (ns <= 0 || ns >= ((int64_t)1 << 32)) {
    set_error("unexpected number of row slices %lld", (long long)ns);
    return fail(CIP_EINVAL);
  }
  // --- slices in (row, channel) order ---------------------------------------------------
  // workspace: s_row (8) + s_idx (24) + keys 2x8 + flags 2x8, s_ch0/ch1 (4+4), ids 2x4
  TTRY(ctx->twork.runs.ensure((size_t)ns * 32));
  TTRY(ctx->twork.run_keys.ensure((size_t)ns * 8));
  TTRY(ctx->twork.run_keys_alt.ensure((size_t)ns * 8));
  TTRY(ctx->twork.run_idx.ensure((size_t)ns * 4));
  TTRY(ctx->twork.run_idx_alt.ensure((size_t)ns * 4));
  TTRY(ctx->twork.run_len.ensure((size_t)ns * 8));
  TTRY(ctx->twork.flags.ensure((size_t)ns * 16));
  TTRY(ctx->scan_out.ensure(sizeof(Range)));
  int64_t* s_row = ctx->twork.runs.as<int64_t>();
  long long* s_idx = reinterpret_cast<long long*>(s_row + ns);
  int32_t* s_ch0 = ctx->twork.run_len.as<int32_t>();
  int32_t* s_ch1 = s_ch0 + ns;
  Range* d_range = ctx->scan_out.as<Range>();
  Range h_range;
  for (int k = 0; k < 3; ++k) { h_range.lo[k] = LLONG_MAX; h_range.hi[k] = LLONG_MIN; }
  TCUDA(cudaMemcpyAsync(d_range, &h_range, sizeof(Range), cudaMemcpyHostToDevice, st));
  TCUDA(cudaStreamSynchronize(st));
  tile_runs_kernel<1><<<rblocks, 256, 0, st>>>(d_uvw, d_freq, nrow, nchan, tile_size[0], tile_size[1], tile_size[2],
                                               nullptr, row_off, s_row, s_ch0, s_idx, d_range);
  TCUDA(cudaMemcpyAsync(&h_range, d_range, sizeof(Range), cudaMemcpyDeviceToHost, st));
  TCUDA(cudaStreamSynchronize(st));
  unsigned long long ext[3];
  int total_bits = 0;
  for (int k = 0; k < 3; ++k) {
    ext[k] = (unsigned long long)(h_range.hi[k] - h_range.lo[k]) + 1ull;
    total_bits += bits_for(ext[k]);
  }
  if (total_bits > 64) {
    set_error("tile index range too large to sort (%d bits)", total_bits);
    return fail(CIP_EINVAL);
  }
  // pad extents to powers of two so the packed key never overflows
  const unsigned long long rv = 1ull << bits_for(ext[1]), rw = 1ull << bits_for(ext[2]);
  uint64_t* keys = ctx->twork.run_keys.as<uint64_t>();
  uint64_t* keys_s = ctx->twork.run_keys_alt.as<uint64_t>();
  uint32_t* ids = ctx->twork.run_idx.as<uint32_t>();
  uint32_t* ids_s = ctx->twork.run_idx_alt.as<uint32_t>();
  slice_finish_kernel<<<nblk(ns, 256), 256, 0, st>>>(s_row, s_ch0, s_idx, ns, (int32_t)nchan, h_range, rv, rw, s_ch1,
                                                     keys, ids);
  TCUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, keys, keys_s, ids, ids_s, ns, 0, total_bits, st));
  TTRY(ctx->twork.cub_tmp.ensure(tmp));
  tmp = ctx->twork.cub_tmp.cap;
  TCUDA(cub::DeviceRadixSort::SortPairs(ctx->twork.cub_tmp.p, tmp, keys, keys_s, ids, ids_s, ns, 0, total_bits, st));
  // --- tiles ----------------------------------------------------------------------------
  int64_t* flag = ctx->twork.flags.as<int64_t>();
  int64_t* tile_of = flag + ns;
  tile_flags_kernel<<<nblk(ns, 256), 256, 0, st>>>(keys_s, ns, flag);
  TCUDA(cub::DeviceScan::InclusiveSum(nullptr, tmp, flag, tile_of, ns, st));
  TTRY(ctx->twork.cub_tmp.ensure(tmp));
  tmp = ctx->twork.cub_tmp.cap;
  TCUDA(cub::DeviceScan::InclusiveSum(ctx->twork.cub_tmp.p, tmp, flag, tile_of, ns, st));
  TCUDA(cudaMemcpyAsync(ctx->scan_host, tile_of + ns - 1, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  TCUDA(cudaStreamSynchronize(st));
  const int64_t nt = *static_cast<int64_t*>(ctx->scan_host);
  // per-tile scratch: first_pos (8) | sizes (8, nt+1) | first_id 2x4 | tile_id 2x4
  TTRY(ctx->twork.group_start.ensure((size_t)(nt + 1) * 40));
  int64_t* first_pos = ctx->twork.group_start.as<int64_t>();
  int64_t* sizes = first_pos + (nt + 1);
  uint32_t* first_id = reinterpret_cast<uint32_t*>(sizes + (nt + 1));
  uint32_t* first_id_s = first_id + (nt + 1);
  uint32_t* tile_id = first_id_s + (nt + 1);
  uint32_t* tile_id_s = tile_id + (nt + 1);
  tile_first_kernel<<<nblk(ns, 256), 256, 0, st>>>(flag, tile_of, ids_s, ns, first_pos, first_id, tile_id);
  TCUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, first_id, first_id_s, tile_id, tile_id_s, nt, 0, 32, st));
  TTRY(ctx->twork.cub_tmp.ensure(tmp));
  tmp = ctx->twork.cub_tmp.cap;
  TCUDA(cub::DeviceRadixSort::SortPairs(ctx->twork.cub_tmp.p, tmp, first_id, first_id_s, tile_id, tile_id_s, nt, 0, 32, st));
  TTRY(R->tile_coords.ensure(sizeof(int64_t) * 3 * nt));
  TTRY(R->tile_offsets.ensure(sizeof(int64_t) * (nt + 1)));
  TTRY(R->slice_irow.ensure(sizeof(int64_t) * ns));
  TTRY(R->slice_ch0.ensure(sizeof(int32_t) * ns));
  TTRY(R->slice_ch1.ensure(sizeof(int32_t) * ns));
  tile_order_kernel<<<nblk(nt + 1, 256), 256, 0, st>>>(tile_id_s, first_pos, ids_s, s_idx, nt, ns, sizes,
                                                       R->tile_coords.as<int64_t>());
  TCUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, sizes, R->tile_offsets.as<int64_t>(), nt + 1, st));
  TTRY(ctx->twork.cub_tmp.ensure(tmp));
  tmp = ctx->twork.cub_tmp.cap;
  TCUDA(cub::DeviceScan::ExclusiveSum(ctx->twork.cub_tmp.p, tmp, sizes, R->tile_offsets.as<int64_t>(), nt + 1, st));
  tile_emit_kernel<<<(unsigned)nt, 128, 0, st>>>(tile_id_s, first_pos, R->tile_offsets.as<int64_t>(), ids_s, s_row,
                                                 s_ch0, s_ch1, row_offset, R->slice_irow.as<int64_t>(),
                                                 R->slice_ch0.as<int32_t>(), R->slice_ch1.as<int32_t>());
  ctx->launches += 12;
  TCUDA(cudaGetLastError());
  TCUDA(cudaStreamSynchronize(st));
#undef TTRY
#undef TCUDA
  R->n_tiles = nt;
  R->n_slices = ns;
  if (n_tiles) *n_tiles = nt;
  if (n_slices) *n_slices = ns;
  *result = R;
  return CIP_OK;
}

extern "C" cip_status cip_tile_fetch(cip_ctx* ctx, const cip_tile_result* r, int64_t* tile_coords,
                                     int64_t* tile_offsets, int64_t* slice_irow, int32_t* slice_ch_start,
                                     int32_t* slice_ch_stop) {
  CIP_REQUIRE(ctx && r, "NULL argument");
  CIP_ENTER_DEVICE(ctx);
  cudaStream_t st = ctx->stream;
  if (tile_offsets)
    CIP_CHECK_CUDA(cudaMemcpyAsync(tile_offsets, r->tile_offsets.p, sizeof(int64_t) * (r->n_tiles + 1), cudaMemcpyDeviceToHost, st));
  if (r->n_tiles && tile_coords)
    CIP_CHECK_CUDA(cudaMemcpyAsync(tile_coords, r->tile_coords.p, sizeof(int64_t) * 3 * r->n_tiles, cudaMemcpyDeviceToHost, st));
  if (r->n_slices) {
    if (slice_irow)
      CIP_CHECK_CUDA(cudaMemcpyAsync(slice_irow, r->slice_irow.p, sizeof(int64_t) * r->n_slices, cudaMemcpyDeviceToHost, st));
    if (slice_ch_start)
      CIP_CHECK_CUDA(cudaMemcpyAsync(slice_ch_start, r->slice_ch0.p, sizeof(int32_t) * r->n_slices, cudaMemcpyDeviceToHost, st));
    if (slice_ch_stop)
      CIP_CHECK_CUDA(cudaMemcpyAsync(slice_ch_stop, r->slice_ch1.p, sizeof(int32_t) * r->n_slices, cudaMemcpyDeviceToHost, st));
  }
  CIP_CHECK_CUDA(cudaStreamSynchronize(st));
  return CIP_OK;
}

extern "C" cip_status cip_tile_gather(cip_ctx* ctx, const cip_tile_result* r, const void* vis, const double* uvw,
                                      int64_t nrow, int64_t nchan, int memspace, void* vis_out, int64_t* vis_offsets,
                                      double* uvw_out) {
  CIP_REQUIRE(ctx && r && vis_out && vis_offsets && uvw_out, "NULL argument");
  CIP_REQUIRE(nrow == r->nrow && nchan == r->nchan, "shape differs from the binned uvw (%lld x %lld)",
              (long long)r->nrow, (long long)r->nchan);
  CIP_REQUIRE(memspace == 0 || memspace == 1, "memspace must be 0 or 1");
  CIP_ENTER_DEVICE(ctx);
  cudaStream_t st = ctx->stream;
  const int64_t ns = r->n_slices, nvis = nrow * nchan;
  if (ns == 0) {
    if (memspace == 0) vis_offsets[0] = 0; else CIP_CHECK_CUDA(cudaMemsetAsync(vis_offsets, 0, 8, st));
    CIP_CHECK_CUDA(cudaStreamSynchronize(st));
    return CIP_OK;
  }
  CIP_REQUIRE(vis && uvw, "NULL argument");
  const float2* d_vis = static_cast<const float2*>(vis);
  const double* d_uvw = uvw;
  float2* d_vout = static_cast<float2*>(vis_out);
  int64_t* d_off = vis_offsets;
  double* d_uout = uvw_out;
  if (memspace == 0) {
    // inputs | outputs packed into one scratch buffer
    size_t need = (size_t)nvis * 8 * 2 + (size_t)nrow * 24 + (size_t)ns * 24 + (size_t)(ns + 1) * 8 + 64;
    CIP_TRY(ctx->t_vis.ensure(need));
    char* p = ctx->t_vis.as<char>();
    d_vis = reinterpret_cast<float2*>(p); p += (size_t)nvis * 8;
    d_vout = reinterpret_cast<float2*>(p); p += (size_t)nvis * 8;
    d_uvw = reinterpret_cast<double*>(p); p += (size_t)nrow * 24;
    d_uout = reinterpret_cast<double*>(p); p += (size_t)ns * 24;
    d_off = reinterpret_cast<int64_t*>(p);
    CIP_CHECK_CUDA(cudaMemcpyAsync((void*)d_vis, vis, (size_t)nvis * 8, cudaMemcpyHostToDevice, st));
    CIP_CHECK_CUDA(cudaMemcpyAsync((void*)d_uvw, uvw, (size_t)nrow * 24, cudaMemcpyHostToDevice, st));
  }
  CIP_TRY(ctx->twork.flags.ensure((size_t)(ns + 1) * 8));
  int64_t* len = ctx->twork.flags.as<int64_t>();
  slice_len_kernel<<<nblk(ns + 1, 256), 256, 0, st>>>(r->slice_ch0.as<int32_t>(), r->slice_ch1.as<int32_t>(), ns, len);
  size_t tmp = 0;
  CIP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, len, d_off, ns + 1, st));
  CIP_TRY(ctx->twork.cub_tmp.ensure(tmp));
  tmp = ctx->twork.cub_tmp.cap;
  CIP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(ctx->twork.cub_tmp.p, tmp, len, d_off, ns + 1, st));
  gather_kernel<<<nblk(ns * 32, 256), 256, 0, st>>>(d_vis, d_uvw, nchan, r->row_offset, r->slice_irow.as<int64_t>(),
                                                    r->slice_ch0.as<int32_t>(), d_off, ns, d_vout, d_uout);
  ctx->launches += 3;
  CIP_CHECK_CUDA(cudaGetLastError());
  if (memspace == 0) {
    CIP_CHECK_CUDA(cudaMemcpyAsync(vis_out, d_vout, (size_t)nvis * 8, cudaMemcpyDeviceToHost, st));
    CIP_CHECK_CUDA(cudaMemcpyAsync(vis_offsets, d_off, (size_t)(ns + 1) * 8, cudaMemcpyDeviceToHost, st));
    CIP_CHECK_CUDA(cudaMemcpyAsync(uvw_out, d_uout, (size_t)ns * 24, cudaMemcpyDeviceToHost, st));
  }
  CIP_CHECK_CUDA(cudaStreamSynchronize(st));
  return CIP_OK;
}
