// Kernel 1 (gridder side): binning of (row, channel) visibilities into uv tiles and w-plane
// windows -- stage G1 of SURVEY.md section 8a, the work ducc0.wgridder.ms2dirty does internally
// before gridding (reference call site src/ska_sdp_cip/invert.py:170-183).
//
// Per row the channels are run-length encoded by their (window, tile_u, tile_v) key (one warp per
// row, ballot on "key differs from the previous channel"); the runs are ordered by key with a
// stable LSD radix sort, so inside a bin they stay in (row, channel) order and the accumulation
// order of the gridder is reproducible; the sorted runs are then expanded into 24-byte
// gridder-ready records (all float64 coordinate arithmetic happens here, once) and cut into work
// items of at most kItemVis records.
#include <algorithm>
#include <cstdlib>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "geom.cuh"

namespace cip {

namespace {

// Sort keys are stored as 32-bit words when they fit (they do up to ~ 8192^2-cell grids): the radix sort moves
// a third fewer bytes per pass.
__device__ __forceinline__ void store_key(void* keys, int64_t pos, uint64_t key, bool key32) {
  if (key32) static_cast<uint32_t*>(keys)[pos] = (uint32_t)key;
  else static_cast<uint64_t*>(keys)[pos] = key;
}
__device__ __forceinline__ uint64_t load_key(const void* keys, int64_t pos, bool key32) {
  return key32 ? (uint64_t)static_cast<const uint32_t*>(keys)[pos] : static_cast<const uint64_t*>(keys)[pos];
}

// One warp per row, lanes stride the channels; a run is a maximal stretch of channels of one row with
// the same (window, tile) bin AND the same register-window lattice position inside the tile, so that
// after the sort all records that share a register window of the gridder's consumer are contiguous.
//   PASS 0: count the runs of each row.
//   PASS 1: write the runs (input order) at row_off[row] and their sort keys.
//   PASS 2: after the sort, expand the samples into gridder-ready records at their sorted positions
//           (outpos[r] = first record of input-order run r); all float64 coordinate arithmetic of the
//           gridder happens here.  Reads are row-contiguous, writes contiguous per run.
template <int PASS, typename R, bool PACK = false>
__global__ void __launch_bounds__(256, 6) runs_kernel(DevInputs in, GridGeom g, int64_t* __restrict__ row_nruns,
                                                   const int64_t* __restrict__ row_off, Run* __restrict__ runs,
                                                   void* __restrict__ keys, const int64_t* __restrict__ outpos,
                                                   Rec<R>* __restrict__ recs, const double* __restrict__ finv, bool key32) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= in.nrow) return;
  const BinConsts kc = bin_consts(g);
  const RowGeom rg = row_geom(g, in.uvw[3 * row], in.uvw[3 * row + 1], in.uvw[3 * row + 2]);
  uint64_t carry = ~0ull;
  int64_t n = 0;
  int carry_ch0 = 0;
  const int64_t base = PASS ? row_off[row] : 0;
  // record pass: the output positions of the row's first 32 runs are fetched up front (one coalesced load,
  // in flight together with the samples) instead of one dependent gather after each ballot
  int64_t pre_out = 0;
  if (PASS == 2 && lane < row_nruns[row]) pre_out = outpos[base + lane];
  for (int64_t c0 = 0; c0 < in.nchan; c0 += 32) {
    const int64_t c = c0 + lane;
    const bool valid = c < in.nchan;
    uint64_t key = ~0ull;
    SampleGeom s;
    BinCell bc;
    // record pass: issue the sample loads before the geometry, which does not depend on them (the pass is
    // bound by memory latency: ncu showed 59 % of the stalls on these loads when they came after the ballot)
    float2 v = make_float2(0.f, 0.f);
    float wgt = 0.f;
    if (PASS == 2 && valid) {
      v = __ldcs(in.vis + row * in.vstride + c);
      wgt = in.wgt ? __ldcs(in.wgt + row * in.wstride + c) : 1.0f;
      if (in.mask && in.mask[row * in.wstride + c] == 0) wgt = 0.0f;
    }
    if (valid) {
      s = sample_geom<PASS == 2>(g, kc, rg, finv[c]);
      bc = bin_cell(g, kc, s);
      key = bin_key(g, bc);
    }
    uint64_t prev = __shfl_up_sync(0xffffffffu, key, 1);
    if (lane == 0) prev = carry;
    const bool start = valid && key != prev;
    const unsigned m = __ballot_sync(0xffffffffu, start);
    if (PASS == 1 && start) {
      const int64_t pos = base + n + __popc(m & ((1u << lane) - 1));
      Run r;
      r.key = key >> kSubKeyBits;
      r.row = (int32_t)row;
      r.ch0 = (int32_t)c;
      runs[pos] = r;
      store_key(keys, pos, key, key32);
    }
    if (PASS == 2) {
      const unsigned upto = m & (0xffffffffu >> (31 - lane));  // starts at or before this lane
      const int64_t rel = n + __popc(upto) - 1;                // index of this sample's run within the row
      const int64_t r = base + rel;                            // input-order index of this sample's run
      int64_t out0 = __shfl_sync(0xffffffffu, pre_out, (int)(rel & 31));
      if (valid && rel >= 32) out0 = outpos[r];
      const int ch0 = upto ? (int)c0 + (31 - __clz(upto)) : carry_ch0;
      if (valid) {
        const bool act = wgt != 0.0f && (v.x != 0.0f || v.y != 0.0f);
        R re = 0, im = 0;
        if (act) {
          re = (R)v.x * (R)wgt;
          im = (R)v.y * (R)wgt;
          if (s.flip) im = -im;
          if (g.do_wstacking && g.nshift != 0.0) {  // V' = V exp(+2 pi i w s); phase reduced in float64
            double ph = s.w * g.nshift;
            ph -= rint(ph);
            R sn, cs;
            if constexpr (sizeof(R) == 4) sincospif(2.0f * (float)ph, &sn, &cs);
            else sincospi(2.0 * ph, &sn, &cs);
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
          *reinterpret_cast<float4*>(reinterpret_cast<RecP*>(recs) + out0 + ((int)c - ch0)) = *reinterpret_cast<const float4*>(&rp);
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
        float4* dst = reinterpret_cast<float4*>(recs + out0 + ((int)c - ch0));
        const float4* src = reinterpret_cast<const float4*>(&rec);
        dst[0] = src[0];
        dst[1] = src[1];
        }
      }
      if (m) carry_ch0 = (int)c0 + (31 - __clz(m));
    }
    n += __popc(m);
    carry = __shfl_sync(0xffffffffu, key, 31);
  }
  if (PASS == 0 && lane == 0) row_nruns[row] = n;
}

// PASS 0 and PASS 1 in one kernel (rows of at most kFusedChan channels): each warp walks its row once, keeps
// the sample keys in shared memory and counts its runs; the CTA reserves a contiguous block of run slots with
// one atomic on a global cursor; the warps then write their runs from the cached keys.  Blocks of runs land in
// CTA arrival order rather than row order, which only permutes the (float) accumulation order inside a bin.
constexpr int kFusedWarps = 8;
constexpr int kFusedChan = 768;
__global__ void __launch_bounds__(32 * kFusedWarps) runs_fused_kernel(DevInputs in, GridGeom g, int64_t* __restrict__ row_nruns,
                                                                      int64_t* __restrict__ row_off, Run* __restrict__ runs,
                                                                      void* __restrict__ keys,
                                                                      const double* __restrict__ finv, int nch32,
                                                                      int64_t cap, bool key32) {
  extern __shared__ uint64_t skeys[];  // [warp][nch32]
  __shared__ int s_cnt[kFusedWarps];
  __shared__ unsigned long long s_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row = (int64_t)blockIdx.x * kFusedWarps + warp;
  const bool active = row < in.nrow;
  uint64_t* mykeys = skeys + (size_t)warp * nch32;
  int n = 0;
  if (active) {
    const BinConsts kc = bin_consts(g);
    const RowGeom rg = row_geom(g, in.uvw[3 * row], in.uvw[3 * row + 1], in.uvw[3 * row + 2]);
    uint64_t carry = ~0ull;
    for (int c0 = 0; c0 < (int)in.nchan; c0 += 32) {
      const int c = c0 + lane;
      const bool valid = c < in.nchan;
      uint64_t key = ~0ull;
      if (valid) {
        const SampleGeom s = sample_geom<false>(g, kc, rg, finv[c]);
        key = bin_key(g, bin_cell(g, kc, s));
      }
      mykeys[c] = key;
      uint64_t prev = __shfl_up_sync(0xffffffffu, key, 1);
      if (lane == 0) prev = carry;
      n += __popc(__ballot_sync(0xffffffffu, valid && key != prev));
      carry = __shfl_sync(0xffffffffu, key, 31);
    }
  }
  if (lane == 0) s_cnt[warp] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int total = 0;
    for (int w = 0; w < kFusedWarps; ++w) total += s_cnt[w];
    s_base = atomicAdd(reinterpret_cast<unsigned long long*>(row_off + in.nrow), (unsigned long long)total);
  }
  __syncthreads();
  if (!active) return;
  int64_t base = (int64_t)s_base;
  for (int w = 0; w < warp; ++w) base += s_cnt[w];
  if (lane == 0) {
    row_off[row] = base;
    row_nruns[row] = n;
  }
  uint64_t carry = ~0ull;
  int m_before = 0;
  for (int c0 = 0; c0 < (int)in.nchan; c0 += 32) {
    const int c = c0 + lane;
    const bool valid = c < in.nchan;
    const uint64_t key = mykeys[c];
    uint64_t prev = __shfl_up_sync(0xffffffffu, key, 1);
    if (lane == 0) prev = carry;
    const bool start = valid && key != prev;
    const unsigned m = __ballot_sync(0xffffffffu, start);
    if (start) {
      const int64_t pos = base + m_before + __popc(m & ((1u << lane) - 1));
      Run r;
      r.key = key >> kSubKeyBits;
      r.row = (int32_t)row;
      r.ch0 = (int32_t)c;
      if (pos < cap) {
        runs[pos] = r;
        store_key(keys, pos, key, key32);
      }
    }
    m_before += __popc(m);
    carry = __shfl_sync(0xffffffffu, key, 31);
  }
}

__global__ void finv_kernel(const double* __restrict__ freq, int64_t nchan, double* __restrict__ finv) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c < nchan) finv[c] = freq[c] / kSpeedOfLight;
}

__global__ void run_len_kernel(const Run* __restrict__ runs, int64_t nruns, int32_t nchan, uint32_t* __restrict__ idx,
                               int32_t* __restrict__ len) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nruns) return;
  Run a = runs[r];
  int32_t stop = nchan;
  if (r + 1 < nruns) {
    Run b = runs[r + 1];
    if (b.row == a.row) stop = b.ch0;
  }
  len[r] = stop - a.ch0;
  idx[r] = (uint32_t)r;
}

// sorted order: gather run + length, flag bin starts
__global__ void gather_kernel(const Run* __restrict__ runs, const int32_t* __restrict__ len,
                              const uint32_t* __restrict__ idx, const void* __restrict__ keys_sorted, bool key32, int64_t nruns,
                              Run* __restrict__ runs_sorted, int64_t* __restrict__ len_sorted,
                              int64_t* __restrict__ start_flag) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nruns) return;
  uint32_t r = idx[i];
  runs_sorted[i] = runs[r];
  len_sorted[i] = len[r];
  start_flag[i] = (i == 0 || (load_key(keys_sorted, i, key32) >> kSubKeyBits) != (load_key(keys_sorted, i - 1, key32) >> kSubKeyBits)) ? 1 : 0;
}

// group g = g-th distinct key; first run of each group, number of items of each group
__global__ void groups_kernel(const int64_t* __restrict__ start_flag, const int64_t* __restrict__ group_of /*inclusive scan*/,
                              const int64_t* __restrict__ visoff, const uint32_t* __restrict__ idx, int64_t nruns,
                              int64_t nrec, int64_t* __restrict__ group_first, int64_t* __restrict__ outpos) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nruns) return;
  outpos[idx[i]] = visoff[i];  // first record of input-order run idx[i]
  if (start_flag[i]) group_first[group_of[i] - 1] = i;
}

__global__ void group_items_kernel(const int64_t* __restrict__ group_first, const int64_t* __restrict__ visoff,
                                   int64_t ngroups, int64_t nruns, int64_t nrec, int64_t* __restrict__ group_nitems) {
  int64_t gidx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gidx > ngroups) return;
  if (gidx == ngroups) { group_nitems[gidx] = 0; return; }
  int64_t b = visoff[group_first[gidx]];
  int64_t e = gidx + 1 < ngroups ? visoff[group_first[gidx + 1]] : nrec;
  group_nitems[gidx] = (e - b + kItemVis - 1) / kItemVis;
}

__global__ void items_kernel(const int64_t* __restrict__ group_first, const int64_t* __restrict__ group_item_off,
                             const int64_t* __restrict__ visoff, const Run* __restrict__ runs_sorted, int64_t ngroups,
                             int64_t nrec, GridGeom g, Item* __restrict__ items, int64_t* __restrict__ win_item_start) {
  int64_t gidx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gidx >= ngroups) return;
  int64_t fr = group_first[gidx];
  int64_t b = visoff[fr];
  int64_t e = gidx + 1 < ngroups ? visoff[group_first[gidx + 1]] : nrec;
  uint64_t key = runs_sorted[fr].key;
  int tv = (int)(key % g.ntv);
  uint64_t q = key / g.ntv;
  int tu = (int)(q % g.ntu);
  int win = (int)(q / g.ntu);
  int64_t it = group_item_off[gidx];
  atomicMin((unsigned long long*)&win_item_start[win], (unsigned long long)it);
  for (int64_t r = b; r < e; r += kItemVis, ++it) {
    Item item;
    item.rec_begin = r;
    item.nrec = (int32_t)std::min<int64_t>(kItemVis, e - r);
    item.tu = tu;
    item.tv = tv;
    item.win = win;
    item.pad[0] = item.pad[1] = 0;
    items[it] = item;
  }
}

__global__ void fill_i64_kernel(int64_t* p, int64_t n, int64_t v) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

template <typename T>
cip_status d2h_scalar(cip_ctx* ctx, const T* dev, T* out) {
  return fetch_small(ctx, dev, sizeof(T), out);
}

inline int nblk(int64_t n, int per) { return (int)((n + per - 1) / per); }

}  // namespace

cip_status run_binning(cip_ctx* ctx, int slot, const DevInputs& in, const GridGeom& g, int fp64, BinResult* out) {
  cudaStream_t st = ctx->stream;
  BinWork& w = ctx->work[slot];
  const int64_t nrow = in.nrow, nchan = in.nchan;
  out->nruns = out->nrec = out->nitems = 0;
  out->win_item_start.assign(g.nwin + 1, 0);
  if (nrow == 0 || nchan == 0) return CIP_OK;
  const int64_t nrec = nrow * nchan;

  // --- runs per row, offsets -------------------------------------------------------
  CIP_TRY(w.row_nruns.ensure(sizeof(int64_t) * (nrow + 1)));
  CIP_TRY(w.row_off.ensure(sizeof(int64_t) * (nrow + 1)));
  int64_t* row_nruns = w.row_nruns.as<int64_t>();
  int64_t* row_off = w.row_off.as<int64_t>();
  CIP_CHECK_CUDA(cudaMemsetAsync(row_nruns + nrow, 0, sizeof(int64_t), st));
  const int rblocks = nblk(nrow * 32, 256);
  CIP_TRY(w.finv.ensure(sizeof(double) * nchan));
  double* finv = w.finv.as<double>();
  finv_kernel<<<nblk(nchan, 256), 256, 0, st>>>(in.freq, nchan, finv);
  ctx->launches++;
  uint64_t nkeys = (uint64_t)g.nwin * g.ntu * g.ntv;
  int bits = 1;
  while (bits < 64 && (nkeys >> bits)) ++bits;
  bits += kSubKeyBits;
  CIP_REQUIRE(bits <= 64, "bin key too wide (%d bits)", bits);
  const bool key32 = bits <= 32 && !getenv("CIP_BIN_KEY64");
  int64_t nruns = 0;
  size_t tmp_bytes = 0;
  Run* runs = nullptr;
  if (nchan <= kFusedChan && !getenv("CIP_BIN_TWO_PASS")) {
    // fused count + write: the run buffers are sized from a guess (what they already hold, or a mean run
    // length of 4 channels) and the kernel drops writes past the capacity; a low guess costs one re-run
    const int nch32 = (int)((nchan + 31) / 32) * 32;
    const size_t smem = sizeof(uint64_t) * kFusedWarps * nch32;
    CIP_CHECK_CUDA(cudaFuncSetAttribute(runs_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t cap = std::max<int64_t>((int64_t)std::min(w.runs.cap / sizeof(Run), w.run_keys.cap / sizeof(uint64_t)),
                                    nrec / 4 + 1024);
    for (int attempt = 0;; ++attempt) {
      CIP_TRY(w.runs.ensure(sizeof(Run) * cap));
      CIP_TRY(w.run_keys.ensure(sizeof(uint64_t) * cap));
      runs = w.runs.as<Run>();
      CIP_CHECK_CUDA(cudaMemsetAsync(row_off + nrow, 0, sizeof(int64_t), st));
      runs_fused_kernel<<<nblk(nrow, kFusedWarps), 32 * kFusedWarps, smem, st>>>(in, g, row_nruns, row_off, runs,
                                                                                 w.run_keys.p, finv, nch32, cap, key32);
      ctx->launches++;
      CIP_TRY(d2h_scalar(ctx, row_off + nrow, &nruns));
      if (nruns <= cap) break;
      CIP_REQUIRE(attempt == 0, "run count changed between two passes");
      cap = nruns;
    }
    CIP_REQUIRE(nruns > 0 && nruns < ((int64_t)1 << 32), "unexpected number of runs %lld", (long long)nruns);
  } else {
    runs_kernel<0, float><<<rblocks, 256, 0, st>>>(in, g, row_nruns, nullptr, nullptr, nullptr, nullptr, (Rec<float>*)nullptr, finv, key32);
    ctx->launches++;
    CIP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, row_nruns, row_off, nrow + 1, st));
    CIP_TRY(w.cub_tmp.ensure(tmp_bytes));
    tmp_bytes = w.cub_tmp.cap;
    CIP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(w.cub_tmp.p, tmp_bytes, row_nruns, row_off, nrow + 1, st));
    ctx->launches++;
    CIP_TRY(d2h_scalar(ctx, row_off + nrow, &nruns));
    CIP_REQUIRE(nruns > 0 && nruns < ((int64_t)1 << 32), "unexpected number of runs %lld", (long long)nruns);
    CIP_TRY(w.runs.ensure(sizeof(Run) * nruns));
    CIP_TRY(w.run_keys.ensure(sizeof(uint64_t) * nruns));
    runs = w.runs.as<Run>();
    runs_kernel<1, float><<<rblocks, 256, 0, st>>>(in, g, nullptr, row_off, runs, w.run_keys.p, nullptr,
                                                   (Rec<float>*)nullptr, finv, key32);
    ctx->launches++;
  }
  CIP_TRY(w.run_keys_alt.ensure(sizeof(uint64_t) * nruns));
  CIP_TRY(w.run_idx.ensure(sizeof(uint32_t) * nruns));
  CIP_TRY(w.run_idx_alt.ensure(sizeof(uint32_t) * nruns));
  CIP_TRY(w.run_len.ensure(sizeof(int32_t) * nruns));
  run_len_kernel<<<nblk(nruns, 256), 256, 0, st>>>(runs, nruns, (int32_t)nchan, w.run_idx.as<uint32_t>(),
                                                   w.run_len.as<int32_t>());
  ctx->launches++;

  // --- stable sort of the runs by key ----------------------------------------------
  if (key32) {
    CIP_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, w.run_keys.as<uint32_t>(),
                                                   w.run_keys_alt.as<uint32_t>(), w.run_idx.as<uint32_t>(),
                                                   w.run_idx_alt.as<uint32_t>(), nruns, 0, bits, st));
    CIP_TRY(w.cub_tmp.ensure(tmp_bytes));
    tmp_bytes = w.cub_tmp.cap;
    CIP_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(w.cub_tmp.p, tmp_bytes, w.run_keys.as<uint32_t>(),
                                                   w.run_keys_alt.as<uint32_t>(), w.run_idx.as<uint32_t>(),
                                                   w.run_idx_alt.as<uint32_t>(), nruns, 0, bits, st));
  } else {
    CIP_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, w.run_keys.as<uint64_t>(),
                                                   w.run_keys_alt.as<uint64_t>(), w.run_idx.as<uint32_t>(),
                                                   w.run_idx_alt.as<uint32_t>(), nruns, 0, bits, st));
    CIP_TRY(w.cub_tmp.ensure(tmp_bytes));
    tmp_bytes = w.cub_tmp.cap;
    CIP_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(w.cub_tmp.p, tmp_bytes, w.run_keys.as<uint64_t>(),
                                                   w.run_keys_alt.as<uint64_t>(), w.run_idx.as<uint32_t>(),
                                                   w.run_idx_alt.as<uint32_t>(), nruns, 0, bits, st));
  }
  ctx->launches += 2 + (bits + 7) / 8;

  // --- sorted runs, record offsets, bins -------------------------------------------
  CIP_TRY(w.runs_sorted.ensure(sizeof(Run) * nruns));
  CIP_TRY(w.run_visoff.ensure(sizeof(int64_t) * (nruns + 1) * 2));  // lengths + offsets
  CIP_TRY(w.flags.ensure(sizeof(int64_t) * nruns * 2));              // start flags + group index
  Run* runs_sorted = w.runs_sorted.as<Run>();
  int64_t* len_sorted = w.run_visoff.as<int64_t>();
  int64_t* visoff = len_sorted + (nruns + 1);
  int64_t* start_flag = w.flags.as<int64_t>();
  int64_t* group_of = start_flag + nruns;
  gather_kernel<<<nblk(nruns, 256), 256, 0, st>>>(runs, w.run_len.as<int32_t>(), w.run_idx_alt.as<uint32_t>(),
                                                  w.run_keys_alt.p, key32, nruns, runs_sorted, len_sorted,
                                                  start_flag);
  ctx->launches++;
  CIP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, len_sorted, visoff, nruns, st));
  CIP_TRY(w.cub_tmp.ensure(tmp_bytes));
  tmp_bytes = w.cub_tmp.cap;
  CIP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(w.cub_tmp.p, tmp_bytes, len_sorted, visoff, nruns, st));
  CIP_CHECK_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, start_flag, group_of, nruns, st));
  CIP_TRY(w.cub_tmp.ensure(tmp_bytes));
  tmp_bytes = w.cub_tmp.cap;
  CIP_CHECK_CUDA(cub::DeviceScan::InclusiveSum(w.cub_tmp.p, tmp_bytes, start_flag, group_of, nruns, st));
  ctx->launches += 2;
  int64_t ngroups = 0;
  CIP_TRY(d2h_scalar(ctx, group_of + nruns - 1, &ngroups));

  CIP_TRY(w.group_start.ensure(sizeof(int64_t) * (ngroups + 1) * 3));
  int64_t* group_first = w.group_start.as<int64_t>();
  int64_t* group_nitems = group_first + (ngroups + 1);
  int64_t* group_item_off = group_nitems + (ngroups + 1);
  CIP_TRY(w.item_idx.ensure(sizeof(int64_t) * nruns));
  int64_t* outpos = w.item_idx.as<int64_t>();
  groups_kernel<<<nblk(nruns, 256), 256, 0, st>>>(start_flag, group_of, visoff, w.run_idx_alt.as<uint32_t>(), nruns,
                                                  nrec, group_first, outpos);
  group_items_kernel<<<nblk(ngroups + 1, 256), 256, 0, st>>>(group_first, visoff, ngroups, nruns, nrec, group_nitems);
  ctx->launches += 2;
  CIP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, group_nitems, group_item_off, ngroups + 1, st));
  CIP_TRY(w.cub_tmp.ensure(tmp_bytes));
  tmp_bytes = w.cub_tmp.cap;
  CIP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(w.cub_tmp.p, tmp_bytes, group_nitems, group_item_off, ngroups + 1, st));
  ctx->launches++;
  int64_t nitems = 0;
  CIP_TRY(d2h_scalar(ctx, group_item_off + ngroups, &nitems));

  CIP_TRY(w.items.ensure(sizeof(Item) * std::max<int64_t>(nitems, 1)));
  CIP_TRY(w.win_item_start.ensure(sizeof(int64_t) * (g.nwin + 1)));
  int64_t* wis = w.win_item_start.as<int64_t>();
  fill_i64_kernel<<<nblk(g.nwin + 1, 256), 256, 0, st>>>(wis, g.nwin + 1, nitems);
  items_kernel<<<nblk(ngroups, 256), 256, 0, st>>>(group_first, group_item_off, visoff, runs_sorted, ngroups, nrec, g,
                                                   w.items.as<Item>(), wis);
  ctx->launches += 2;

  // --- records ---------------------------------------------------------------------
  CIP_TRY(w.recs.ensure(recs_bytes(nrec)));
  if (fp64)
    runs_kernel<2, double><<<rblocks, 256, 0, st>>>(in, g, row_nruns, row_off, nullptr, nullptr, outpos,
                                                    w.recs.as<Rec<double>>(), finv, key32);
  else if (g.W <= kFastSupport)
    runs_kernel<2, float, true><<<rblocks, 256, 0, st>>>(in, g, row_nruns, row_off, nullptr, nullptr, outpos,
                                                         w.recs.as<Rec<float>>(), finv, key32);
  else
    runs_kernel<2, float><<<rblocks, 256, 0, st>>>(in, g, row_nruns, row_off, nullptr, nullptr, outpos,
                                                   w.recs.as<Rec<float>>(), finv, key32);
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());

  std::vector<int64_t> h(g.nwin + 1);
  if (sizeof(int64_t) * (g.nwin + 1) <= kSmallHostBytes) {
    CIP_TRY(fetch_small(ctx, wis, sizeof(int64_t) * (g.nwin + 1), h.data()));
  } else {
    CIP_CHECK_CUDA(cudaMemcpyAsync(h.data(), wis, sizeof(int64_t) * (g.nwin + 1), cudaMemcpyDeviceToHost, st));
    CIP_CHECK_CUDA(cudaStreamSynchronize(st));
  }
  h[g.nwin] = nitems;
  for (int w = g.nwin - 1; w >= 0; --w) h[w] = std::min(h[w], h[w + 1]);  // empty windows start where the next does
  out->win_item_start = h;
  out->nruns = nruns;
  out->nrec = nrec;
  out->nitems = nitems;
  return CIP_OK;
}

}  // namespace cip
