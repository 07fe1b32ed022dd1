// Kernel 2: w-stacked gridding with the exponential-of-semicircle kernel -- stage G2 of
// SURVEY.md section 8a, the hot loop of ducc0.wgridder.ms2dirty (reference call site
// src/ska_sdp_cip/invert.py:170-183).
//
// Design ("residue owner"): one CTA per work item = a slice of the records of one uv tile
// (kTile x kTile support origins) and one window of kWin w planes.  The tile's sub-grid
// (kSub x kSub cells x kWin planes) lives in shared memory.  Warp 0 (producer) streams the
// 24-byte records with coalesced loads, evaluates the ES kernel separably (8 u-, 8 v-, 8 w-
// weights per visibility) and hands batches of 32 visibilities to warp 1 (consumer) through a
// two-stage shared-memory ring guarded by named barriers.  In the consumer, lane (a, b) owns the
// grid cells with u index = a (mod 8) and v index = b or b+4 (mod 8): the support of ANY
// visibility (W <= 8) holds exactly one cell of each residue class, so every lane always has
// exactly one cell per class to update, no two lanes ever touch the same cell, and no atomics
// or shuffles are needed.  The lane keeps its 2 cells x 8 planes in registers (packed f32x2
// FMAs: FFMA2) while consecutive visibilities fall on the same cells -- along a run of channels
// they move by a fraction of a cell -- and folds them into the shared sub-grid (plain,
// conflict-free 128-bit read-modify-write; the row pitch kSub is odd) only when the support
// origin moves.  The finished sub-grid is added to the global plane stack with vector
// reductions (red.global.add.v4.f32), skipping empty cells.
#include "geom.cuh"

namespace cip {

namespace {

constexpr int kStages = 2;
constexpr int kGridThreads = 64;

template <typename R> struct Cplx;
template <> struct Cplx<float> { using type = float2; using cell4 = float4; };
template <> struct Cplx<double> { using type = double2; };

template <typename R>
struct alignas(16) Stage {
  float ku[kBatch][8];    // weight of u residue class r for visibility k (0 outside the support)
  float kv[kBatch][8];    // v classes, stored as pairs: [2*b + s] = class b + 4*s
  float kw[kBatch][16];   // the 8 plane weights, each duplicated (operand of the packed FMA)
  typename Cplx<R>::type val[kBatch];
  uint32_t info[kBatch];  // bits 0-4 i0, 5-9 j0, bit 31: support origin differs from the previous visibility
};

__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ float2 unpack2(unsigned long long v) {
  float2 r;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
  return r;
}
// d = a * b + d on two packed floats (FFMA2 on sm_100a)
__device__ __forceinline__ void ffma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
__device__ __forceinline__ unsigned long long fmul2(unsigned long long a, unsigned long long b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// exp(beta_w (sqrt(t) - 1)) with the hardware approximations; bw2 = beta_w * log2(e)
__device__ __forceinline__ float es_fast(float x, float bw2) {
  float t = fmaf(-x, x, 1.0f);
  float s;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(fmaxf(t, 0.0f)));
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fmaf(bw2, s, -bw2)));
  return t >= 0.0f ? e : 0.0f;
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// ---- producer -------------------------------------------------------------------------
template <typename R, typename REC>
__device__ __forceinline__ void produce(Stage<R>& st, const REC& rec, bool valid, int lane, uint32_t& carry,
                                        const GridGeom& g, float bw2, float two_over_w) {
  const int W = g.W;
  uint32_t cxy = valid ? (rec.cell & 0x3ffu) : carry;
  uint32_t prev = __shfl_up_sync(0xffffffffu, cxy, 1);
  if (lane == 0) prev = carry;
  carry = __shfl_sync(0xffffffffu, cxy, 31);
  const int i0 = cxy & 31, j0 = (cxy >> 5) & 31, p0 = (rec.cell >> 10) & 15;
  float ku[8], kv[8], kw[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    int da = (r - i0) & 7, db = (r - j0) & 7, dc = r - p0;
    float xu = fmaf(rec.fu + (float)da, two_over_w, -1.0f);
    float xv = fmaf(rec.fv + (float)db, two_over_w, -1.0f);
    float xw = fmaf(rec.fw + (float)dc, two_over_w, -1.0f);
    ku[r] = da < W ? es_fast(xu, bw2) : 0.0f;
    kv[r] = db < W ? es_fast(xv, bw2) : 0.0f;
    kw[r] = (dc >= 0 && dc < W) ? es_fast(xw, bw2) : 0.0f;
  }
  if (!g.do_wstacking) {
#pragma unroll
    for (int r = 0; r < 8; ++r) kw[r] = r == 0 ? 1.0f : 0.0f;
  }
  float4* pu = reinterpret_cast<float4*>(st.ku[lane]);
  pu[0] = make_float4(ku[0], ku[1], ku[2], ku[3]);
  pu[1] = make_float4(ku[4], ku[5], ku[6], ku[7]);
  float4* pv = reinterpret_cast<float4*>(st.kv[lane]);
  pv[0] = make_float4(kv[0], kv[4], kv[1], kv[5]);
  pv[1] = make_float4(kv[2], kv[6], kv[3], kv[7]);
  float4* pw = reinterpret_cast<float4*>(st.kw[lane]);
  pw[0] = make_float4(kw[0], kw[0], kw[1], kw[1]);
  pw[1] = make_float4(kw[2], kw[2], kw[3], kw[3]);
  pw[2] = make_float4(kw[4], kw[4], kw[5], kw[5]);
  pw[3] = make_float4(kw[6], kw[6], kw[7], kw[7]);
  typename Cplx<R>::type v;
  v.x = valid ? (R)rec.re : (R)0;
  v.y = valid ? (R)rec.im : (R)0;
  st.val[lane] = v;
  st.info[lane] = cxy | (cxy != prev ? 0x80000000u : 0u);
}

// ---- consumer -------------------------------------------------------------------------
template <typename R> struct Acc;

template <>
struct Acc<float> {
  unsigned long long a[2][8];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
      for (int c = 0; c < 8; ++c) a[s][c] = 0ull;
  }
  // sub-grid cell = float4 {re(c), im(c), re(c+1), im(c+1)} for plane pair pp = c/2
  __device__ __forceinline__ void fold(float4* sg, int i, int j0s, int j1s) {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      int j = s ? j1s : j0s;
#pragma unroll
      for (int pp = 0; pp < 4; ++pp) {
        float4* p = sg + ((pp * kSub + i) * kSub + j);
        float4 v = *p;
        float2 lo = unpack2(a[s][2 * pp]), hi = unpack2(a[s][2 * pp + 1]);
        v.x += lo.x; v.y += lo.y; v.z += hi.x; v.w += hi.y;
        *p = v;
        a[s][2 * pp] = 0ull;
        a[s][2 * pp + 1] = 0ull;
      }
    }
  }
  __device__ __forceinline__ void update(const Stage<float>& st, int k, int a8, int b4) {
    float ku = st.ku[k][a8];
    float2 kv = *reinterpret_cast<const float2*>(&st.kv[k][2 * b4]);
    float2 val = st.val[k];
    float w0 = ku * kv.x, w1 = ku * kv.y;
    unsigned long long t0 = pack2(val.x * w0, val.y * w0), t1 = pack2(val.x * w1, val.y * w1);
    const ulonglong2* kw = reinterpret_cast<const ulonglong2*>(st.kw[k]);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      ulonglong2 w = kw[q];
      ffma2(a[0][2 * q], t0, w.x);
      ffma2(a[0][2 * q + 1], t0, w.y);
      ffma2(a[1][2 * q], t1, w.x);
      ffma2(a[1][2 * q + 1], t1, w.y);
    }
  }
};

template <>
struct Acc<double> {
  double2 a[2][8];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
      for (int c = 0; c < 8; ++c) a[s][c] = make_double2(0.0, 0.0);
  }
  // sub-grid cell = double2 per plane: layout [plane][i][j]
  __device__ __forceinline__ void fold(double2* sg, int i, int j0s, int j1s) {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      int j = s ? j1s : j0s;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        double2* p = sg + ((c * kSub + i) * kSub + j);
        double2 v = *p;
        v.x += a[s][c].x;
        v.y += a[s][c].y;
        *p = v;
        a[s][c] = make_double2(0.0, 0.0);
      }
    }
  }
  __device__ __forceinline__ void update(const Stage<double>& st, int k, int a8, int b4) {
    float ku = st.ku[k][a8];
    float2 kv = *reinterpret_cast<const float2*>(&st.kv[k][2 * b4]);
    double2 val = st.val[k];
    double w0 = (double)ku * (double)kv.x, w1 = (double)ku * (double)kv.y;
    double tr0 = val.x * w0, ti0 = val.y * w0, tr1 = val.x * w1, ti1 = val.y * w1;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      double w = (double)st.kw[k][2 * c];
      a[0][c].x = fma(tr0, w, a[0][c].x);
      a[0][c].y = fma(ti0, w, a[0][c].y);
      a[1][c].x = fma(tr1, w, a[1][c].x);
      a[1][c].y = fma(ti1, w, a[1][c].y);
    }
  }
};

template <typename R> struct SubGrid;
template <> struct SubGrid<float> { using cell = float4; static constexpr int ncell = 4 * kSub * kSub; };
template <> struct SubGrid<double> { using cell = double2; static constexpr int ncell = 8 * kSub * kSub; };

// ---- the kernel -----------------------------------------------------------------------
template <typename R, typename REC>
__global__ void __launch_bounds__(kGridThreads) grid_kernel(const Item* __restrict__ items, int64_t item_base,
                                                            const REC* __restrict__ recs, GridGeom g,
                                                            R* __restrict__ grid, int plane_base, int nslots) {
  using Cell = typename SubGrid<R>::cell;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Cell* sg = reinterpret_cast<Cell*>(smem_raw);
  Stage<R>* stages = reinterpret_cast<Stage<R>*>(smem_raw + sizeof(Cell) * SubGrid<R>::ncell);

  const Item item = items[item_base + blockIdx.x];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nrec = item.nrec;
  const int nb = (nrec + kBatch - 1) / kBatch;

  {  // zero the sub-grid
    Cell z;
    memset(&z, 0, sizeof(z));
    for (int t = tid; t < SubGrid<R>::ncell; t += kGridThreads) sg[t] = z;
  }
  __syncthreads();

  if (warp == 0) {
    // ---------------- producer ----------------
    const float bw2 = g.beta_w * 1.4426950408889634f;
    const float two_over_w = 2.0f / (float)g.W;
    const REC* src = recs + item.rec_begin;
    uint32_t carry = 0xffffffffu;
    REC cur;
    bool cur_valid = lane < nrec;
    if (cur_valid) cur = src[lane]; else memset(&cur, 0, sizeof(cur));
    for (int b = 0; b < nb; ++b) {
      const int s = b & 1;
      REC nxt;
      int kn = (b + 1) * kBatch + lane;
      bool nxt_valid = kn < nrec;
      if (nxt_valid) nxt = src[kn]; else memset(&nxt, 0, sizeof(nxt));
      if (b >= kStages) bar_sync(3 + s, kGridThreads);  // wait until the consumer released stage s
      produce<R, REC>(stages[s], cur, cur_valid, lane, carry, g, bw2, two_over_w);
      __threadfence_block();
      bar_arrive(1 + s, kGridThreads);
      cur = nxt;
      cur_valid = nxt_valid;
    }
  } else {
    // ---------------- consumer ----------------
    const int a8 = lane & 7, b4 = lane >> 3;
    Acc<R> acc;
    acc.clear();
    int ci = a8, cj0 = b4, cj1 = b4 + 4;  // current cells of this lane (sub-grid coordinates)
    for (int b = 0; b < nb; ++b) {
      const int s = b & 1;
      bar_sync(1 + s, kGridThreads);
      const Stage<R>& st = stages[s];
      const int n = min(kBatch, nrec - b * kBatch);
#pragma unroll 4
      for (int k = 0; k < n; ++k) {
        const uint32_t info = st.info[k];
        if (info & 0x80000000u) {  // warp-uniform: the support origin moved
          acc.fold(sg, ci, cj0, cj1);
          const int i0 = info & 31, j0 = (info >> 5) & 31;
          ci = i0 + ((a8 - i0) & 7);
          cj0 = j0 + ((b4 - j0) & 7);
          cj1 = j0 + ((b4 + 4 - j0) & 7);
        }
        acc.update(st, k, a8, b4);
      }
      if (b + kStages < nb) bar_arrive(3 + s, kGridThreads);
    }
    acc.fold(sg, ci, cj0, cj1);
  }
  __syncthreads();

  // ---------------- sub-grid -> global plane stack ----------------
  const int P0 = item.win * g.win_step;
  const int su0 = item.tu * kTile - kHalo + g.nu / 2;  // shifted -> FFT order: (s + nu/2) mod nu
  const int sv0 = item.tv * kTile - kHalo + g.nv / 2;
  if constexpr (sizeof(R) == 4) {
    // pairs of v-adjacent cells whose global index is even, so the 16-byte reduction is aligned
    constexpr int kPairs = (kSub + 1) / 2;
    const int odd = sv0 & 1;
    for (int t = tid; t < 4 * kSub * kPairs; t += kGridThreads) {
      int jp = t % kPairs, q = t / kPairs;
      int i = q % kSub, pp = q / kSub;
      int j = 2 * jp - odd;
      const bool ok0 = j >= 0, ok1 = j + 1 < kSub;
      float4 c0 = ok0 ? sg[(pp * kSub + i) * kSub + j] : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 c1 = ok1 ? sg[(pp * kSub + i) * kSub + j + 1] : make_float4(0.f, 0.f, 0.f, 0.f);
      int gi = (su0 + i) % g.nu;
      int gj = (sv0 + j + g.nv) % g.nv;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        int plane = P0 + 2 * pp + h;
        if (plane >= g.nplanes) continue;
        int slot = plane - plane_base;
        if (slot < 0 || slot >= nslots) continue;
        float a = h ? c0.z : c0.x, bb = h ? c0.w : c0.y, c = h ? c1.z : c1.x, d = h ? c1.w : c1.y;
        if (a == 0.f && bb == 0.f && c == 0.f && d == 0.f) continue;
        float* dst = grid + 2 * (((int64_t)slot * g.nu + gi) * g.nv + gj);
        if (ok0 && ok1) {
          red_add_v4(dst, a, bb, c, d);
        } else if (ok0) {
          atomicAdd(dst, a);
          atomicAdd(dst + 1, bb);
        } else {
          atomicAdd(dst + 2, c);
          atomicAdd(dst + 3, d);
        }
      }
    }
  } else {
    for (int t = tid; t < 8 * kSub * kSub; t += kGridThreads) {
      int j = t % kSub, q = t / kSub;
      int i = q % kSub, c = q / kSub;
      int plane = P0 + c;
      int slot = plane - plane_base;
      if (plane >= g.nplanes || slot < 0 || slot >= nslots) continue;
      double2 v = reinterpret_cast<double2*>(sg)[(c * kSub + i) * kSub + j];
      if (v.x == 0.0 && v.y == 0.0) continue;
      int gi = (su0 + i) % g.nu;
      if (gi < 0) gi += g.nu;
      int gj = (sv0 + j) % g.nv;
      if (gj < 0) gj += g.nv;
      double* dst = reinterpret_cast<double*>(grid) + 2 * (((int64_t)slot * g.nu + gi) * g.nv + gj);
      atomicAdd(dst, v.x);
      atomicAdd(dst + 1, v.y);
    }
  }
}

// Wide supports (W > 8, only reached for very small epsilon): plain scatter with global atomics.
template <typename R, typename REC>
__global__ void __launch_bounds__(128) grid_wide_kernel(const Item* __restrict__ items, int64_t item_base,
                                                        const REC* __restrict__ recs, GridGeom g, R* __restrict__ grid,
                                                        int plane_base, int nslots) {
  const Item item = items[item_base + blockIdx.x];
  const int W = g.W;
  const int su0 = item.tu * kTile - kHalo + g.nu / 2, sv0 = item.tv * kTile - kHalo + g.nv / 2;
  for (int k = threadIdx.x; k < item.nrec; k += blockDim.x) {
    const REC rec = recs[item.rec_begin + k];
    if (rec.re == 0 && rec.im == 0) continue;
    const int i0 = rec.cell & 31, j0 = (rec.cell >> 5) & 31, p0 = rec.cell >> 10;
    float kv[kMaxSupport];
    for (int b = 0; b < W; ++b) kv[b] = es_kernel((2.0f * (rec.fv + b) - W) / W, g.beta_w);
    for (int c = 0; c < (g.do_wstacking ? W : 1); ++c) {
      int plane = p0 + c, slot = plane - plane_base;
      if (plane >= g.nplanes || slot < 0 || slot >= nslots) continue;
      float kw = g.do_wstacking ? es_kernel((2.0f * (rec.fw + c) - W) / W, g.beta_w) : 1.0f;
      for (int a = 0; a < W; ++a) {
        float ku = es_kernel((2.0f * (rec.fu + a) - W) / W, g.beta_w) * kw;
        int gi = (su0 + i0 + a) % g.nu;
        if (gi < 0) gi += g.nu;
        R tr = (R)rec.re * (R)ku, ti = (R)rec.im * (R)ku;
        for (int b = 0; b < W; ++b) {
          int gj = (sv0 + j0 + b) % g.nv;
          if (gj < 0) gj += g.nv;
          R* dst = grid + 2 * (((int64_t)slot * g.nu + gi) * g.nv + gj);
          atomicAdd(dst, tr * (R)kv[b]);
          atomicAdd(dst + 1, ti * (R)kv[b]);
        }
      }
    }
  }
}

template <typename R, typename REC>
cip_status launch(cip_ctx* ctx, const GridGeom& g, int64_t item_begin, int64_t item_end, void* grid, int plane_base,
                  int nslots) {
  const Item* items = ctx->items.as<Item>();
  const REC* recs = ctx->recs.as<REC>();
  const int64_t n = item_end - item_begin;
  if (n <= 0) return CIP_OK;
  CIP_REQUIRE(n < ((int64_t)1 << 31), "too many work items");
  if (g.W <= kFastSupport) {
    size_t smem = sizeof(typename SubGrid<R>::cell) * SubGrid<R>::ncell + sizeof(Stage<R>) * kStages;
    auto kern = grid_kernel<R, REC>;
    CIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)n, kGridThreads, smem, ctx->stream>>>(items, item_begin, recs, g, static_cast<R*>(grid), plane_base,
                                                          nslots);
  } else {
    grid_wide_kernel<R, REC><<<(unsigned)n, 128, 0, ctx->stream>>>(items, item_begin, recs, g, static_cast<R*>(grid),
                                                                   plane_base, nslots);
  }
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

}  // namespace

cip_status run_gridder(cip_ctx* ctx, const GridGeom& g, int fp64, int64_t item_begin, int64_t item_end, void* grid,
                       int plane_base, int nslots) {
  if (fp64) return launch<double, VisRec64>(ctx, g, item_begin, item_end, grid, plane_base, nslots);
  return launch<float, VisRec>(ctx, g, item_begin, item_end, grid, plane_base, nslots);
}

}  // namespace cip
