// Kernel 1 (gridder side): counting sort of (row, channel) visibilities into uv tiles, w-plane windows and
// register-window lattice positions -- stage G1 of SURVEY.md section 8a, the work ducc0.wgridder.ms2dirty does
// internally before gridding (reference call site src/ska_sdp_cip/invert.py:170-183).
//
// The sort key of a sample is dense: (window, tile_u, tile_v, lattice_u, lattice_v); all samples of one key share
// one 8 x 8 x 8 accumulator window of the gridder.  Three passes over the samples' geometry, no sort of anything:
//   1. histogram   one warp per row, lanes stride the channels; a maximal stretch of equal keys inside a
//                  32-channel chunk (a "segment") adds its length to count[key] with one atomic;
//   2. scan        count -> padded k-groups of 8 records and jobs per key, exclusive prefix over the key table
//                  (tile sums, one-CTA scan of the sums, apply); the apply pass leaves the first record position
//                  of every key in the table, writes the jobs (key order = window-major) and the zero records
//                  that pad every key's group to whole k-groups;
//   3. records     one warp per row again: all float64 coordinate arithmetic of the gridder happens here, once;
//                  each segment reserves its positions with one atomic on the key's cursor and writes the
//                  gridder-ready records (16 bytes packed on the float32 W <= 8 path, else 32).
// Nothing is read back by the host: sizes are capacities, the totals and an overflow flag stay on the device
// (mirrored into mapped host memory), and every consumer skips its work when the flag is set; the caller checks it
// once after the whole call has drained, grows the buffers and repeats the call (never in steady state).
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "geom.cuh"

namespace cip {

namespace {

constexpr int kScanTile = 2048;      // table entries per CTA of the scan passes
constexpr int kScanThreads = 256;
constexpr int kScanPer = kScanTile / kScanThreads;
constexpr int64_t kMaxTableEntries = (int64_t)1 << 30;  // 4 GB of u32

__global__ void finv_kernel(const double* __restrict__ freq, int64_t nchan, double* __restrict__ finv) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c < nchan) finv[c] = freq[c] / kSpeedOfLight;
}

// ---- pass 1: histogram of the keys ------------------------------------------------------------------------
__global__ void __launch_bounds__(256) hist_kernel(DevInputs in, GridGeom g, BinConsts kc, const double* __restrict__ finv,
                                                   uint32_t* __restrict__ table) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= in.nrow) return;
  const RowGeom rg = row_geom(g, in.uvw[3 * row], in.uvw[3 * row + 1], in.uvw[3 * row + 2]);
  const RowSpan sp = row_span(in, row);
  for (int c0 = sp.c_begin & ~31; c0 < sp.c_end; c0 += 32) {
    const int c = c0 + lane;
    uint32_t key = kNoKey;
    if (c >= sp.c_begin && c < sp.c_end) {
      const SampleGeom s = sample_geom<false>(g, kc, rg, finv[c]);
      if (!(in.w_hi > 0.0) || (s.w >= in.w_lo && s.w < in.w_hi)) key = bin_key(g, bin_cell(g, kc, s));
    }
    uint32_t prev = __shfl_up_sync(0xffffffffu, key, 1);
    const unsigned bound = __ballot_sync(0xffffffffu, lane == 0 || key != prev);  // segment starts (incl. skipped ones)
    if (((bound >> lane) & 1u) && key != kNoKey) {
      const unsigned after = lane == 31 ? 0u : bound >> (lane + 1);
      const int len = after ? __ffs(after) : 32 - lane;  // up to the next boundary or the end of the chunk
      atomicAdd(table + key, (uint32_t)min(len, sp.c_end - c));
    }
  }
}

// ---- pass 2: scan of the key table ------------------------------------------------------------------------
__device__ __forceinline__ uint2 entry_cost(uint32_t count) {
  const uint32_t ngroups = (count + kGroup - 1) / kGroup;
  return make_uint2(ngroups, (ngroups + kJobGroups - 1) / kJobGroups);
}

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const uint32_t* __restrict__ table, int64_t n,
                                                                   uint2* __restrict__ tile_sums) {
  __shared__ uint2 s_part[kScanThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanPer;
  uint32_t gsum = 0, jsum = 0;
  if (base + kScanPer <= n) {
    const uint4 a = *reinterpret_cast<const uint4*>(table + base), b = *reinterpret_cast<const uint4*>(table + base + 4);
    const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < kScanPer; ++i) {
      const uint2 e = entry_cost(v[i]);
      gsum += e.x;
      jsum += e.y;
    }
  } else {
    for (int i = 0; i < kScanPer; ++i)
      if (base + i < n) {
        const uint2 e = entry_cost(table[base + i]);
        gsum += e.x;
        jsum += e.y;
      }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    gsum += __shfl_xor_sync(0xffffffffu, gsum, o);
    jsum += __shfl_xor_sync(0xffffffffu, jsum, o);
  }
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = make_uint2(gsum, jsum);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint2 t = make_uint2(0, 0);
    for (int w = 0; w < kScanThreads / 32; ++w) {
      t.x += s_part[w].x;
      t.y += s_part[w].y;
    }
    tile_sums[blockIdx.x] = t;
  }
}

// One CTA: exclusive prefix of the tile sums in place, totals and capacity check into the header.
__global__ void __launch_bounds__(1024) scan_tiles_kernel(uint2* __restrict__ tile_sums, int64_t ntiles, BinHeader* __restrict__ hdr,
                                                          uint32_t cap_groups, uint32_t cap_jobs, uint32_t* __restrict__ call_flags,
                                                          uint32_t* __restrict__ win_job_total, BinHeader* __restrict__ hdr_host,
                                                          uint32_t* __restrict__ flags_host) {
  __shared__ uint2 s_warp[32];
  __shared__ uint2 s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = make_uint2(0, 0);
  __syncthreads();
  for (int64_t t0 = 0; t0 < ntiles; t0 += 1024) {
    const int64_t t = t0 + threadIdx.x;
    const uint2 mine = t < ntiles ? tile_sums[t] : make_uint2(0, 0);
    uint2 inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t gx = __shfl_up_sync(0xffffffffu, inc.x, o), jx = __shfl_up_sync(0xffffffffu, inc.y, o);
      if (lane >= o) {
        inc.x += gx;
        inc.y += jx;
      }
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      uint2 w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t gx = __shfl_up_sync(0xffffffffu, w.x, o), jx = __shfl_up_sync(0xffffffffu, w.y, o);
        if (lane >= o) {
          w.x += gx;
          w.y += jx;
        }
      }
      s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const uint2 carry = s_carry;
    const uint2 wbase = warp ? s_warp[warp - 1] : make_uint2(0, 0);
    if (t < ntiles) tile_sums[t] = make_uint2(carry.x + wbase.x + inc.x - mine.x, carry.y + wbase.y + inc.y - mine.y);
    __syncthreads();
    if (threadIdx.x == 0) s_carry = make_uint2(carry.x + s_warp[31].x, carry.y + s_warp[31].y);
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const uint2 tot = s_carry;
    BinHeader h;
    h.ngroups = tot.x;
    h.njobs = tot.y;
    h.overflow = (tot.x > cap_groups ? 1u : 0u) | (tot.y > cap_jobs ? 2u : 0u);
    h.pad = 0;
    *hdr = h;
    *win_job_total = tot.y;
    if (h.overflow) atomicOr(call_flags, h.overflow);
    // mirrors for the host's check after the call has drained: the largest demand seen so far in this call
    hdr_host->ngroups = max(hdr_host->ngroups, tot.x);
    hdr_host->njobs = max(hdr_host->njobs, tot.y);
    if (h.overflow) *flags_host |= h.overflow;
    __threadfence_system();
  }
}

// Apply: per key its first record position into the table, the jobs, the padding records, the first job of
// every window.  REC_BYTES = 16 (packed float32 records) or 32.
template <int REC_BYTES>
__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(uint32_t* __restrict__ table, int64_t n, const uint2* __restrict__ tile_prefix,
                                                                  const BinHeader* __restrict__ hdr, GridGeom g, int64_t entries_per_window,
                                                                  Job* __restrict__ jobs, unsigned char* __restrict__ recs,
                                                                  uint32_t* __restrict__ win_job_start) {
  __shared__ uint2 s_warp[kScanThreads / 32];
  if (hdr->overflow) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanPer;
  uint32_t cnt[kScanPer];
  uint2 cost[kScanPer];
  uint2 sum = make_uint2(0, 0);
#pragma unroll
  for (int i = 0; i < kScanPer; ++i) {
    cnt[i] = base + i < n ? table[base + i] : 0u;
    cost[i] = entry_cost(cnt[i]);
    sum.x += cost[i].x;
    sum.y += cost[i].y;
  }
  uint2 inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t gx = __shfl_up_sync(0xffffffffu, inc.x, o), jx = __shfl_up_sync(0xffffffffu, inc.y, o);
    if (lane >= o) {
      inc.x += gx;
      inc.y += jx;
    }
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint2 off = tile_prefix[blockIdx.x];
  for (int w = 0; w < warp; ++w) {
    off.x += s_warp[w].x;
    off.y += s_warp[w].y;
  }
  off.x += inc.x - sum.x;
  off.y += inc.y - sum.y;
  const uint32_t snap = g.W <= kFastSupport ? 9 - g.W : 1;
#pragma unroll
  for (int i = 0; i < kScanPer; ++i) {
    const int64_t idx = base + i;
    if (idx >= n) break;
    if (idx % entries_per_window == 0) win_job_start[idx / entries_per_window] = off.y;
    if (cnt[i]) {
      const uint32_t ngroups = cost[i].x;
      table[idx] = off.x * kGroup;  // the key's record cursor
      const KeyParts kp = key_parts(g, (uint32_t)idx);
      for (uint32_t j = 0; j < cost[i].y; ++j) {
        Job job;
        job.group_begin = off.x + j * kJobGroups;
        job.ngroups = min((uint32_t)kJobGroups, ngroups - j * kJobGroups);
        job.tu = (uint16_t)kp.tu;
        job.tv = (uint16_t)kp.tv;
        job.win = (uint16_t)kp.win;
        job.lu = (uint8_t)kp.lu;
        job.lv = (uint8_t)kp.lv;
        jobs[off.y + j] = job;
      }
      // zero records up to the end of the last k-group; the packed ones carry the window's lattice origin so that
      // a consumer that tracks the window by the records' origins sees no change
      const uint32_t lv = (uint32_t)kp.lv, lu = (uint32_t)kp.lu;
      const uint4 zero = make_uint4(REC_BYTES == 16 ? (lu * snap) | ((lv * snap) << 3) : 0u, 0u, 0u, 0u);
      for (uint32_t r = cnt[i]; r < ngroups * kGroup; ++r) {
        uint4* dst = reinterpret_cast<uint4*>(recs + ((size_t)off.x * kGroup + r) * REC_BYTES);
        dst[0] = zero;
        if (REC_BYTES == 32) dst[1] = make_uint4(0u, 0u, 0u, 0u);
      }
      off.x += ngroups;
      off.y += cost[i].y;
    }
  }
}

// Work items of the register-window gridders (grid.cu): a job starts an item when it is the first of its (window,
// tile) bin or crosses a multiple of kItemGroups k-groups; the item runs up to the next such job.  The start flag
// goes into bit 31 of the job's ngroups, the starts into a list (in job order inside a CTA of this kernel, CTAs in
// arrival order, so neighbouring items mostly stay neighbours and share grid lines in L2).
__global__ void __launch_bounds__(256) items_kernel(Job* __restrict__ jobs, const BinHeader* __restrict__ hdr,
                                                    uint32_t* __restrict__ items, uint32_t* __restrict__ nitems) {
  __shared__ uint32_t s_warp[8];
  __shared__ uint32_t s_base;
  if (hdr->overflow) return;
  const uint32_t njobs = hdr->njobs;
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (blockIdx.x * blockDim.x >= njobs) return;
  bool start = false;
  if (j < njobs) {
    const Job me = jobs[j];
    start = j == 0;
    if (j > 0) {
      const Job prev = jobs[j - 1];
      start = prev.tu != me.tu || prev.tv != me.tv || prev.win != me.win ||
              (prev.group_begin / kItemGroups) != (me.group_begin / kItemGroups);
    }
    if (start) jobs[j].ngroups = me.ngroups | kJobStartsItem;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned m = __ballot_sync(0xffffffffu, start);
  if (lane == 0) s_warp[warp] = __popc(m);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t tot = 0;
    for (int w = 0; w < 8; ++w) {
      const uint32_t c = s_warp[w];
      s_warp[w] = tot;
      tot += c;
    }
    s_base = atomicAdd(nitems, tot);
  }
  __syncthreads();
  if (start) items[s_base + s_warp[warp] + __popc(m & ((1u << lane) - 1))] = j;
}

// ---- pass 3: records --------------------------------------------------------------------------------------
// One warp per row, lanes stride the channels.  Reads are row-contiguous; a segment's records are contiguous.
template <typename R, bool PACK, bool WIDE>
__global__ void __launch_bounds__(256, 6) records_kernel(DevInputs in, GridGeom g, BinConsts kc, const double* __restrict__ finv,
                                                         uint32_t* __restrict__ table, const BinHeader* __restrict__ hdr,
                                                         Rec<R>* __restrict__ recs) {
  if (hdr->overflow) return;
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= in.nrow) return;
  const RowGeom rg = row_geom(g, in.uvw[3 * row], in.uvw[3 * row + 1], in.uvw[3 * row + 2]);
  const RowSpan sp = row_span(in, row);
  for (int c0 = sp.c_begin & ~31; c0 < sp.c_end; c0 += 32) {
    const int c = c0 + lane;
    const bool valid = c >= sp.c_begin && c < sp.c_end;
    // the sample loads are issued before the geometry, which does not depend on them (the pass is bound by
    // memory latency)
    using V = typename std::conditional<WIDE, double, float>::type;  // precision of the inputs
    using V2 = typename std::conditional<WIDE, double2, float2>::type;
    V2 v;
    v.x = v.y = 0;
    V wgt = 0;
    if (valid) {
      v = __ldcs(reinterpret_cast<const V2*>(in.vis) + sp.vbase + c);
      wgt = in.wgt ? __ldcs(reinterpret_cast<const V*>(in.wgt) + sp.wbase + c) : (V)1;
      if (in.mask && in.mask[sp.wbase + c] == 0) wgt = 0;
    }
    uint32_t key = kNoKey;
    SampleGeom s;
    BinCell bc;
    if (valid) {
      s = sample_geom<true>(g, kc, rg, finv[c]);
      bc = bin_cell(g, kc, s);
      if (!(in.w_hi > 0.0) || (s.w >= in.w_lo && s.w < in.w_hi)) key = bin_key(g, bc);
    }
    const uint32_t prev = __shfl_up_sync(0xffffffffu, key, 1);
    const unsigned bound = __ballot_sync(0xffffffffu, lane == 0 || key != prev);
    const int lead = 31 - __clz(bound & (0xffffffffu >> (31 - lane)));  // first lane of this lane's segment
    uint32_t pos = 0;
    if (lead == lane && key != kNoKey) {
      const unsigned after = lane == 31 ? 0u : bound >> (lane + 1);
      const int len = after ? __ffs(after) : 32 - lane;
      pos = atomicAdd(table + key, (uint32_t)min(len, sp.c_end - c));
    }
    pos = __shfl_sync(0xffffffffu, pos, lead) + (uint32_t)(lane - lead);
    if (key == kNoKey) continue;
    const bool act = wgt != 0 && (v.x != 0 || v.y != 0);
    R re = 0, im = 0;
    if (act) {
      // the product in the wider of the two precisions (float32 inputs: exactly the float32 product of before)
      using P = typename std::conditional<(sizeof(V) > sizeof(R)), V, R>::type;
      re = (R)((P)v.x * (P)wgt);
      im = (R)((P)v.y * (P)wgt);
      if (s.flip) im = -im;
      if (g.do_wstacking && g.nshift != 0.0) {  // V' = V exp(+2 pi i w s); phase reduced in float64
        const double ph = s.w * g.nshift;
        R sn, cs;
        if constexpr (sizeof(R) == 4) {
          sincos_turns(ph, &sn, &cs);
        } else {
          sincospi(2.0 * (ph - rint_magic(ph)), &sn, &cs);
        }
        const R t = re * cs - im * sn;
        im = re * sn + im * cs;
        re = t;
      }
    }
    if constexpr (PACK) {  // float32 register-window gridder: 16-byte record, one 128-bit store
      RecP rp;
      rp.geo = (uint64_t)(bc.i0r | (bc.j0r << 3) | (bc.p0r << 6)) | (to_fix18(s.dfu) << 10) | (to_fix18(s.dfv) << 28) |
               (to_fix18(s.dfw) << 46);
      rp.re = (float)re;
      rp.im = (float)im;
      *reinterpret_cast<float4*>(reinterpret_cast<RecP*>(recs) + pos) = *reinterpret_cast<const float4*>(&rp);
    } else {
      Rec<R> rec;
      rec.cell = bc.i0r | (bc.j0r << 5) | (bc.p0r << 10);
      // bits 16-20 / 21-25: origin of the gridder's register window (i0r, j0r snapped to the lattice);
      // wide supports have no windows and p0r may need the high bits
      if (kc.fast) rec.cell |= ((bc.lu * kc.snap) << 16) | ((bc.lv * kc.snap) << 21);
      if constexpr (sizeof(R) == 4) {
        rec.fu = s.fu;
        rec.fv = s.fv;
        rec.fw = s.fw;
      } else {
        rec.fu = to_fix32(s.dfu);
        rec.fv = to_fix32(s.dfv);
        rec.fw = to_fix32(s.dfw);
      }
      rec.re = re;
      rec.im = im;
      if constexpr (sizeof(R) == 4) rec.pad[0] = rec.pad[1] = 0.f;
      float4* dst = reinterpret_cast<float4*>(recs + pos);
      const float4* src = reinterpret_cast<const float4*>(&rec);
      dst[0] = src[0];
      dst[1] = src[1];
    }
  }
}

inline int nblk(int64_t n, int per) { return (int)((n + per - 1) / per); }

int64_t entries_per_window(const GridGeom& g) { return (int64_t)g.ntu * g.ntv * g.nlat * g.nlat; }

}  // namespace

int max_windows_per_pass(const GridGeom& g) {
  int64_t lim = kMaxTableEntries;
  if (const char* e = getenv("CIP_BIN_TABLE_ENTRIES")) lim = std::max<int64_t>(atoll(e), 1);  // tests: force window chunks
  return (int)std::max<int64_t>(1, std::min<int64_t>(g.nwin, lim / std::max<int64_t>(entries_per_window(g), 1)));
}

cip_status run_binning(cip_ctx* ctx, int slot, const DevInputs& in, const GridGeom& g, int fp64, bool want_items) {
  NvtxRange nvtx("cip:binning");
  cudaStream_t st = ctx->stream;
  BinWork& w = ctx->work[slot];
  const int64_t nrow = in.nrow, nchan = in.nchan;
  const int nwin_pass = g.win_end - g.win_begin;
  CIP_REQUIRE(nwin_pass >= 1, "empty window range");
  const int64_t epw = entries_per_window(g);
  const int64_t n = epw * nwin_pass;
  CIP_REQUIRE(epw <= kMaxTableEntries && n < ((int64_t)1 << 32), "key table too large (%lld entries per window)", (long long)epw);
  CIP_REQUIRE(g.ntu < 65536 && g.ntv < 65536 && g.nwin < 65536, "grid or plane stack too large for the job format");
  const int64_t nrec = in.voff ? in.nvis : nrow * nchan;
  CIP_REQUIRE(nrec < ((int64_t)1 << 32) - ((int64_t)1 << 28), "too many samples for one binning pass");
  const bool pack = !fp64 && g.W <= kFastSupport;
  const size_t rec_bytes = pack ? 16 : 32;

  // capacities: what the buffers hold already, or a first guess (12.5 % padding, one job per 64 records); a pass
  // that needs more sets the overflow flag and the call is repeated with the demand it reported
  const int64_t want_groups = std::max<int64_t>(w.cap_groups, (nrec + nrec / 8) / kGroup + 1024);
  const int64_t want_jobs = std::max<int64_t>(w.cap_jobs, nrec / 64 + 1024);
  CIP_TRY(w.recs.ensure((size_t)want_groups * kGroup * 32));  // sized for 32-byte records: both formats fit
  CIP_TRY(w.jobs.ensure(sizeof(Job) * (size_t)want_jobs));
  w.cap_groups = (int64_t)(w.recs.cap / (kGroup * 32));
  w.cap_jobs = (int64_t)(w.jobs.cap / sizeof(Job));
  (void)rec_bytes;
  CIP_TRY(w.table.ensure(sizeof(uint32_t) * (size_t)n + 64));
  const int64_t ntiles = (n + kScanTile - 1) / kScanTile;
  CIP_TRY(w.tile_sums.ensure(sizeof(uint2) * (size_t)ntiles));
  CIP_TRY(w.hdr.ensure(sizeof(BinHeader)));
  CIP_TRY(w.win_job_start.ensure(sizeof(uint32_t) * (size_t)(g.nwin + 2)));
  CIP_TRY(w.finv.ensure(sizeof(double) * std::max<int64_t>(nchan, 1)));
  uint32_t* table = w.table.as<uint32_t>();
  double* finv = w.finv.as<double>();
  BinHeader* hdr = w.hdr.as<BinHeader>();
  uint32_t* wjs = w.win_job_start.as<uint32_t>();

  CIP_CHECK_CUDA(cudaMemsetAsync(table, 0, sizeof(uint32_t) * (size_t)n, st));
  if (nrec == 0) {
    CIP_CHECK_CUDA(cudaMemsetAsync(hdr, 0, sizeof(BinHeader), st));
    CIP_CHECK_CUDA(cudaMemsetAsync(wjs, 0, sizeof(uint32_t) * (size_t)(g.nwin + 2), st));
    return CIP_OK;
  }
  finv_kernel<<<nblk(nchan, 256), 256, 0, st>>>(in.freq, nchan, finv);
  const int rblocks = nblk(nrow * 32, 256);
  const BinConsts kc = bin_consts(g);
  // One window and no |w| slab: every sample's key has window 0 whatever its plane, so the key pass skips the
  // w coordinate (the record pass computes it for the records; both passes agree on the keys).
  GridGeom gk = g;
  if (g.nwin == 1 && !(in.w_hi > 0.0)) gk.do_wstacking = 0;
  hist_kernel<<<rblocks, 256, 0, st>>>(in, gk, kc, finv, table);
  scan_reduce_kernel<<<(unsigned)ntiles, kScanThreads, 0, st>>>(table, n, w.tile_sums.as<uint2>());
  scan_tiles_kernel<<<1, 1024, 0, st>>>(w.tile_sums.as<uint2>(), ntiles, hdr, (uint32_t)std::min<int64_t>(w.cap_groups, 0xffffffffll),
                                        (uint32_t)std::min<int64_t>(w.cap_jobs, 0xffffffffll), ctx->call_flags.as<uint32_t>(),
                                        wjs + nwin_pass, static_cast<BinHeader*>(ctx->flags_host_dev) + 1,
                                        static_cast<uint32_t*>(ctx->flags_host_dev));
  if (pack) {
    scan_apply_kernel<16><<<(unsigned)ntiles, kScanThreads, 0, st>>>(table, n, w.tile_sums.as<uint2>(), hdr, g, epw, w.jobs.as<Job>(),
                                                                     w.recs.as<unsigned char>(), wjs);
    if (in.wide) records_kernel<float, true, true><<<rblocks, 256, 0, st>>>(in, g, kc, finv, table, hdr, w.recs.as<Rec<float>>());
    else records_kernel<float, true, false><<<rblocks, 256, 0, st>>>(in, g, kc, finv, table, hdr, w.recs.as<Rec<float>>());
  } else {
    scan_apply_kernel<32><<<(unsigned)ntiles, kScanThreads, 0, st>>>(table, n, w.tile_sums.as<uint2>(), hdr, g, epw, w.jobs.as<Job>(),
                                                                     w.recs.as<unsigned char>(), wjs);
    if (fp64 && in.wide) records_kernel<double, false, true><<<rblocks, 256, 0, st>>>(in, g, kc, finv, table, hdr, w.recs.as<Rec<double>>());
    else if (fp64) records_kernel<double, false, false><<<rblocks, 256, 0, st>>>(in, g, kc, finv, table, hdr, w.recs.as<Rec<double>>());
    else if (in.wide) records_kernel<float, false, true><<<rblocks, 256, 0, st>>>(in, g, kc, finv, table, hdr, w.recs.as<Rec<float>>());
    else records_kernel<float, false, false><<<rblocks, 256, 0, st>>>(in, g, kc, finv, table, hdr, w.recs.as<Rec<float>>());
  }
  ctx->launches += 6;
  if (want_items) {
    // capacity: one item per job at most
    CIP_TRY(w.items.ensure(sizeof(uint32_t) * (size_t)w.cap_jobs + 64));
    uint32_t* nitems = w.items.as<uint32_t>() + w.cap_jobs;  // [0] item count, [1] the gridder's work counter
    CIP_CHECK_CUDA(cudaMemsetAsync(nitems, 0, 2 * sizeof(uint32_t), st));
    items_kernel<<<nblk(w.cap_jobs, 256), 256, 0, st>>>(w.jobs.as<Job>(), hdr, w.items.as<uint32_t>(), nitems);
    ctx->launches++;
  }
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

cip_status binning_overflowed(cip_ctx* ctx, bool* again) {
  // mirrors written by scan_tiles_kernel: [0] flags, BinHeader at +1 (largest demand of the call)
  volatile uint32_t* flags = static_cast<volatile uint32_t*>(ctx->flags_host);
  const BinHeader* demand = static_cast<const BinHeader*>(ctx->flags_host) + 1;
  *again = flags[0] != 0;
  if (*again) {
    for (BinWork& w : ctx->work) {
      const int64_t g = (int64_t)demand->ngroups + demand->ngroups / 16 + 1024, j = (int64_t)demand->njobs + demand->njobs / 16 + 1024;
      w.cap_groups = std::max(w.cap_groups, g);
      w.cap_jobs = std::max(w.cap_jobs, j);
    }
  }
  return CIP_OK;
}

}  // namespace cip
