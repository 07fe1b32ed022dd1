// Kernel 2: w-stacked gridding with the exponential-of-semicircle kernel -- stage G2 of
// SURVEY.md section 8a, the hot loop of ducc0.wgridder.ms2dirty (reference call site
// src/ska_sdp_cip/invert.py:170-183).
//
// float32 grids with W <= 8 are gridded by the tcgen05 kernel of grid_tc.cu.  This file holds the other paths, all
// working on the jobs of bin.cu (the records of one uv tile, one window of kWin w planes and one register-window
// lattice position), each CTA striding over the job list:
//
// * float32, grid_warp_kernel ("rank-1 tiles"): one warp per item.  Per batch of 32 visibilities the warp
//   first evaluates the ES kernel separably (one visibility per lane) and stages the 8 x 8 products
//   ku kv and the 8 complex plane values V kw in shared memory, then updates the 8 x 8 x 8 register
//   window: groups of 8 visibilities that share it go through the tensor cores as C[64 x 16] += A[64 x 8]
//   B[8 x 16] (mma.sync m16n8k8, every operand split into two TF32 terms, three products per tile, tiles
//   added in float32 -- accuracy of the FFMA path, see tests/accuracy_modes.py); the group in which the
//   window moves takes the rank-1 updates with 16 FFMA2 per lane and visibility.  Consecutive visibilities
//   share the window (its origin is snapped to a lattice of 9-W cells and the records are sorted by
//   lattice position); when the window moves it is added to the global plane stack with vector reductions.
//   CIP_GRID_MMA=0 selects FFMA2 for all groups, =2 accumulates inside the tensor core (faster by 0.2 ms,
//   but its truncating accumulator costs 4x in error on dense tiles at tight kernels).
//
// * float64, grid_kernel<double> ("residue owner"): a producer warp evaluates the kernel in float64 and
//   hands batches to a consumer warp whose lane (a, b) owns the cells u = a, v = b or b+4 (mod 8) of the
//   window; windows are folded into a sub-grid in shared memory that is flushed at the end of the item.
#include <cstdlib>

#include "geom.cuh"

namespace cip {

namespace {

constexpr int kStages = 2;
constexpr int kGridThreads = 64;

using u64 = unsigned long long;

// A job as the kernels below see it.
struct Item {
  int64_t rec_begin;
  int32_t nrec, tu, tv, win;
};
__device__ __forceinline__ Item job_item(const Job& job, const GridGeom& g) {
  (void)g;
  Item it;
  it.rec_begin = (int64_t)job.group_begin * kGroup;
  it.nrec = (int32_t)(job.ngroups & ~kJobStartsItem) * kGroup;
  it.tu = job.tu;
  it.tv = job.tv;
  it.win = job.win;
  return it;
}
// The jobs of windows [win_first, win_last] (bin.cu leaves the first job of every window of the pass on the device).
struct JobList {
  const Job* jobs;
  const uint32_t* win_job_start;
  const BinHeader* hdr;
  int win_first, win_last;
};
__device__ __forceinline__ void job_range(const JobList& jl, const GridGeom& g, uint32_t* jb, uint32_t* je) {
  if (jl.hdr->overflow) {
    *jb = *je = 0;
    return;
  }
  *jb = jl.win_job_start[jl.win_first - g.win_begin];
  *je = jl.win_job_start[jl.win_last + 1 - g.win_begin];
}

// ---- small PTX helpers ----------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// Rendezvous of the producer and the consumer warp (named barrier 1; barrier 0 is __syncthreads).
// A blocked warp costs no issue slots, unlike a polling mbarrier wait.
__device__ __forceinline__ void pair_sync() { asm volatile("bar.sync 1, 64;" ::: "memory"); }

__device__ __forceinline__ u64 pack2(float lo, float hi) {
  u64 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ float2 unpack2(u64 v) {
  float2 r;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
  return r;
}
// d = a * b + d on two packed floats (FFMA2 on sm_100a)
__device__ __forceinline__ void ffma2(u64& d, u64 a, u64 b) { asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b)); }
// packed float32 pairs: one issue slot for two operations (the FMA pipe is busy for two cycles)
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 fsub2(u64 a, u64 b) { u64 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ void unpack2u(u64 v, uint32_t& lo, uint32_t& hi) { asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// ---- producer -> consumer hand-over ------------------------------------------------------
template <typename R> struct StageVal;
template <> struct StageVal<double> { using type = double2; };  // (re, im)

template <typename R>
struct Stage;

// float64 path: weights in float64 (the float32 hardware approximations of es_fast would cap the
// accuracy near 1e-6), no pair duplication (the consumer uses scalar DFMA)
template <>
struct alignas(16) Stage<double> {
  double ku[kBatch][8];   // weight of u residue class r
  double kv[kBatch][8];   // weight of v residue class r
  double kw[kBatch][8];   // the 8 plane weights of the window
  double2 val[kBatch];    // (re, im)
  uint32_t org[kBatch];
  uint32_t moved;
  uint32_t pad[3];
};

template <typename R>
__device__ __forceinline__ Rec<R> load_rec(const Rec<R>* __restrict__ recs, int64_t k, bool valid) {
  Rec<R> r;
  float4* d = reinterpret_cast<float4*>(&r);
  const float4* p = reinterpret_cast<const float4*>(recs + k);
  d[0] = valid ? __ldg(p) : make_float4(0.f, 0.f, 0.f, 0.f);
  d[1] = valid ? __ldg(p + 1) : make_float4(0.f, 0.f, 0.f, 0.f);
  return r;
}

__device__ __forceinline__ RecP load_recp(const RecP* __restrict__ recs, int64_t k, bool valid) {
  RecP r;
  *reinterpret_cast<float4*>(&r) = valid ? __ldg(reinterpret_cast<const float4*>(recs + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
  return r;
}

// Register-window origin of a record and the "window moved" mask of the batch (common to both precisions).
__device__ __forceinline__ uint32_t window_origin(uint32_t cell, bool valid, int lane, uint32_t& carry, int snap,
                                                  uint32_t* moved) {
  // register window: origin snapped to a lattice of `snap` cells, so that nearby supports share it
  // (computed by the record pass of the binning, bits 16-25 of `cell`)
  (void)snap;
  const uint32_t wxy = (cell >> 16) & 0x3ffu;
  const uint32_t cxy = valid ? wxy : carry;
  uint32_t prev = __shfl_up_sync(0xffffffffu, cxy, 1);
  if (lane == 0) prev = carry;
  carry = __shfl_sync(0xffffffffu, cxy, 31);
  *moved = __ballot_sync(0xffffffffu, cxy != prev);
  return cxy;
}

__device__ __forceinline__ void produce(Stage<double>& st, const Rec<double>& rec, bool valid, int lane, uint32_t& carry,
                                        const GridGeom& g, float, float, int snap) {
  const int W = g.W;
  const int i0 = rec.cell & 31, j0 = (rec.cell >> 5) & 31, p0 = (rec.cell >> 10) & 15;
  uint32_t moved;
  const uint32_t cxy = window_origin(rec.cell, valid, lane, carry, snap, &moved);
  const double fu = from_fix32(rec.fu), fv = from_fix32(rec.fv), fw = from_fix32(rec.fw);
  const double two_over_w = 2.0 / (double)W, bw = g.beta_wd;
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const int da = (r - i0) & 7, db = (r - j0) & 7, dc = r - p0;
    st.ku[lane][r] = da < W ? es_kernel(fma(fu + (double)da, two_over_w, -1.0), bw) : 0.0;
    st.kv[lane][r] = db < W ? es_kernel(fma(fv + (double)db, two_over_w, -1.0), bw) : 0.0;
    double kw = (dc >= 0 && dc < W) ? es_kernel(fma(fw + (double)dc, two_over_w, -1.0), bw) : 0.0;
    if (!g.do_wstacking) kw = r == 0 ? 1.0 : 0.0;
    st.kw[lane][r] = kw;
  }
  st.val[lane] = valid ? make_double2(rec.re, rec.im) : make_double2(0.0, 0.0);
  st.org[lane] = cxy;
  if (lane == 0) st.moved = moved;
}

// ---- consumer -------------------------------------------------------------------------
template <typename R> struct SubGrid;
template <> struct SubGrid<double> { using cell = double2; static constexpr int ncell = 8 * kSub * kSub; };

template <typename R> struct Acc;

template <>
struct Acc<double> {
  struct Ops {
    double ku, kv0, kv1;
    double2 val;
    double kw[8];
  };
  double2 a[2][8];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
      for (int c = 0; c < 8; ++c) a[s][c] = make_double2(0.0, 0.0);
  }
  struct Group {
    const Stage<double>* st;
    int k0, a8, b4;
  };
  static __device__ __forceinline__ Group group(const Stage<double>& st, int k0, int a8, int b4) {
    return Group{&st, k0, a8, b4};
  }
  template <int U>
  static __device__ __forceinline__ Ops load(const Group& gp) {
    const Stage<double>& st = *gp.st;
    const int k = min(gp.k0 + U, kBatch - 1);
    Ops o;
    o.ku = st.ku[k][gp.a8];
    o.kv0 = st.kv[k][gp.b4];
    o.kv1 = st.kv[k][gp.b4 + 4];
    o.val = st.val[k];
#pragma unroll
    for (int c = 0; c < 8; ++c) o.kw[c] = st.kw[k][c];
    return o;
  }
  __device__ __forceinline__ void update(const Ops& o) {
    const double w0 = o.ku * o.kv0, w1 = o.ku * o.kv1;
    const double tr0 = o.val.x * w0, ti0 = o.val.y * w0, tr1 = o.val.x * w1, ti1 = o.val.y * w1;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const double w = o.kw[c];
      a[0][c].x = fma(tr0, w, a[0][c].x);
      a[0][c].y = fma(ti0, w, a[0][c].y);
      a[1][c].x = fma(tr1, w, a[1][c].x);
      a[1][c].y = fma(ti1, w, a[1][c].y);
    }
  }
  // sub-grid cell = double2 per plane: layout [plane][i][j]
  __device__ __forceinline__ void fold(double2* sg, int i, int j0s, int j1s) {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      const int j = s ? j1s : j0s;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        double2* p = sg + ((c * kSub + i) * kSub + j);
        double2 v = *p;
        v.x += a[s][c].x;
        v.y += a[s][c].y;
        *p = v;
      }
    }
    clear();
  }
};

template <typename R>
constexpr size_t grid_smem_bytes() {
  return sizeof(typename SubGrid<R>::cell) * SubGrid<R>::ncell + sizeof(Stage<R>) * kStages;
}

// One visibility of a group of eight (compile-time position U): prefetch the operands of the next
// visibility into `nxt`, fold if the register window moved (warp-uniform; only in CHECK groups),
// accumulate `cur`.  Every stage always holds kBatch entries (the producer pads with zero
// visibilities), so there is no bounds check and the two operand sets simply alternate roles.
template <typename R, int U, bool CHECK>
__device__ __forceinline__ void consume1(const Stage<R>& st, const typename Acc<R>::Group& gp, int k0, uint32_t moved,
                                         const typename Acc<R>::Ops& cur, typename Acc<R>::Ops& nxt, Acc<R>& acc,
                                         typename SubGrid<R>::cell* sg, int& ci, int& cj0, int& cj1, int a8, int b4) {
  nxt = Acc<R>::template load<U + 1>(gp);
  if (CHECK && ((moved >> (k0 + U)) & 1u)) {
    acc.fold(sg, ci, cj0, cj1);
    const uint32_t o = st.org[k0 + U];
    const int wi = o & 31, wj = (o >> 5) & 31;
    ci = wi + ((a8 - wi) & 7);
    cj0 = wj + ((b4 - wj) & 7);
    cj1 = wj + ((b4 + 4 - wj) & 7);
  }
  acc.update(cur);
}

template <typename R, bool CHECK>
__device__ __forceinline__ void consume8(const Stage<R>& st, const typename Acc<R>::Group& gp, int k0, uint32_t moved,
                                         typename Acc<R>::Ops& x, Acc<R>& acc, typename SubGrid<R>::cell* sg, int& ci,
                                         int& cj0, int& cj1, int a8, int b4) {
  typename Acc<R>::Ops y;
  consume1<R, 0, CHECK>(st, gp, k0, moved, x, y, acc, sg, ci, cj0, cj1, a8, b4);
  consume1<R, 1, CHECK>(st, gp, k0, moved, y, x, acc, sg, ci, cj0, cj1, a8, b4);
  consume1<R, 2, CHECK>(st, gp, k0, moved, x, y, acc, sg, ci, cj0, cj1, a8, b4);
  consume1<R, 3, CHECK>(st, gp, k0, moved, y, x, acc, sg, ci, cj0, cj1, a8, b4);
  consume1<R, 4, CHECK>(st, gp, k0, moved, x, y, acc, sg, ci, cj0, cj1, a8, b4);
  consume1<R, 5, CHECK>(st, gp, k0, moved, y, x, acc, sg, ci, cj0, cj1, a8, b4);
  consume1<R, 6, CHECK>(st, gp, k0, moved, x, y, acc, sg, ci, cj0, cj1, a8, b4);
  consume1<R, 7, CHECK>(st, gp, k0, moved, y, x, acc, sg, ci, cj0, cj1, a8, b4);
}

// ---- the kernel -----------------------------------------------------------------------
template <typename R>
__global__ void __launch_bounds__(kGridThreads) grid_kernel(JobList jl, const Rec<R>* __restrict__ recs, GridGeom g, R* __restrict__ grid,
                                                            int plane_base, int nslots) {
  using Cell = typename SubGrid<R>::cell;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Cell* sg = reinterpret_cast<Cell*>(smem_raw);
  Stage<R>* stages = reinterpret_cast<Stage<R>*>(smem_raw + sizeof(Cell) * SubGrid<R>::ncell);
  uint32_t jb, je;
  job_range(jl, g, &jb, &je);
  for (uint32_t jix = jb + blockIdx.x; jix < je; jix += gridDim.x) {
  const Item item = job_item(jl.jobs[jix], g);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nrec = item.nrec;
  const int nb = (nrec + kBatch - 1) / kBatch;

  // Which of the two warps consumes: the sub-partition (SMSP) of a warp is its hardware slot modulo 4
  // and a two-warp CTA occupies an aligned pair of slots, so a fixed role would put every consumer of
  // the SM on SMSPs 1 and 3.  Alternating the role with bit 2 of the slot spreads the (FMA-heavy)
  // consumers of co-resident CTAs over all four SMSPs.  Only a heuristic: any choice is correct.
  __shared__ int s_consumer_warp;
  if (tid == 0) {
    uint32_t hw;
    asm volatile("mov.u32 %0, %%warpid;" : "=r"(hw));
    s_consumer_warp = (hw >> 2) & 1;
  }
  {  // zero the sub-grid
    Cell z;
    memset(&z, 0, sizeof(z));
    for (int t = tid; t < SubGrid<R>::ncell; t += kGridThreads) sg[t] = z;
  }
  __syncthreads();

  if (warp != s_consumer_warp) {
    // ---------------- producer ----------------
    const float bw2 = g.beta_w * 1.4426950408889634f;
    const float two_over_w = 2.0f / (float)g.W;
    const int snap = 9 - g.W;  // windows of 8 cells hold the supports of `snap` consecutive origins
    const int64_t base = item.rec_begin;
    uint32_t carry = 0xffffffffu;
    bool cur_valid = lane < nrec;
    Rec<R> cur = load_rec(recs, base + lane, cur_valid);
    for (int b = 0; b < nb; ++b) {
      const int s = b % kStages;
      const int kn = (b + 1) * kBatch + lane;
      const bool nxt_valid = kn < nrec;
      Rec<R> nxt = load_rec(recs, base + kn, nxt_valid);
      produce(stages[s], cur, cur_valid, lane, carry, g, bw2, two_over_w, snap);
      pair_sync();  // rendezvous b: batch b is ready, and the consumer has finished batch b-1 (stage s^1 is free)
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
      const int s = b % kStages;
      pair_sync();  // rendezvous b
      const Stage<R>& st = stages[s];
      const uint32_t moved = st.moved;
      typename Acc<R>::Ops x = Acc<R>::template load<0>(Acc<R>::group(st, 0, a8, b4));
#pragma unroll 1
      for (int k0 = 0; k0 < kBatch; k0 += 8) {
        const typename Acc<R>::Group gp = Acc<R>::group(st, k0, a8, b4);
        // most groups stay inside one register window: branch-free fast path
        if (((moved >> k0) & 0xffu) == 0) consume8<R, false>(st, gp, k0, moved, x, acc, sg, ci, cj0, cj1, a8, b4);
        else consume8<R, true>(st, gp, k0, moved, x, acc, sg, ci, cj0, cj1, a8, b4);
      }
    }
    acc.fold(sg, ci, cj0, cj1);
  }
  __syncthreads();

  // ---------------- sub-grid -> global plane stack ----------------
  const int P0 = item.win * g.win_step;
  const int su0 = item.tu * kTile - kHalo + g.nu / 2;  // shifted -> FFT order: (s + nu/2) mod nu
  const int sv0 = item.tv * kTile - kHalo + g.nv / 2;
  {
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
  __syncthreads();  // the sub-grid and the stages are reused by the CTA's next job
  }
}

// ---- warp-autonomous float32 gridder ("rank-1 tiles") -----------------------------------------
// One warp = one CTA = one work item.  Per batch of 32 visibilities the warp first PRODUCES (one
// visibility per lane): the separable ES weights, the 64 products A[r_u][r_v] = ku[r_u] kv[r_v] of the
// 8 x 8 register window (by residue class, 0 outside the support) and the 8 complex plane values
// B[p] = V kw[p]; then it CONSUMES them: the update of the window is the rank-1 product A (x) B, and
// lane (a, pq) owns the 8 x 2 block {u = a, all 8 v residues} x {planes 2pq, 2pq+1}: two 128-bit and
// two 64-bit shared-memory loads (8 A values, 2 complex B values) feed 16 FFMA2 whose A operand is the
// register-broadcast form (R.F32) -- 6 shared-memory wavefronts per visibility instead of the 14 of
// the two-warp design, which ncu showed to be bound by shared-memory bandwidth.  No hand-over between
// warps, hence no barriers besides __syncwarp.  Sub-grid, fold and epilogue are those of grid_kernel.
constexpr int kARow = 64 * 4 + 16;  // bytes per visibility of the A stage; the pad makes the producer's
constexpr int kBRow = 16 * 4 + 16;  // 128-bit stores (one row per lane) bank-conflict free
// The tensor-path kernels stage the FACTORS ku[8], kv[8] instead of their 64 products (80 instead of 272 bytes per
// visibility; the consumer multiplies the 16 products of its fragment itself -- the FMA pipe has room once the
// window update is on the tensor pipe, the shared-memory pipe does not).  Factor row, in words: ku[0..3] at 0..3,
// kv[0..7] at 4..11, ku[4..7] at 16..19 -- the same bank argument as for the B row.
constexpr int kKRow = 80;
constexpr int warp_stage_bytes(bool fact) { return kBatch * ((fact ? kKRow : kARow) + kBRow + 4); }
constexpr int warp_smem_bytes(bool fact) { return warp_stage_bytes(fact); }

template <int OFF>
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ ulonglong2 lds_u2(uint32_t addr) {
  ulonglong2 v;
  asm volatile("ld.shared.v2.b64 {%0, %1}, [%2+%3];" : "=l"(v.x), "=l"(v.y) : "r"(addr), "n"(OFF));
  return v;
}
__device__ __forceinline__ void sts_f4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void sts_u2(uint32_t addr, u64 a, u64 b) {
  asm volatile("st.shared.v2.b64 [%0], {%1, %2};" ::"r"(addr), "l"(a), "l"(b) : "memory");
}
// ES tap exp(beta_w (sqrt(1 - x^2) - 1)), 0 where 1 - x^2 < 0: the square root of a negative number is NaN, which
// goes through the fma and ex2, and max.f32 returns its other operand for a NaN input -- one FMNMX instead of a
// compare and a select per tap (24 taps per visibility)
__device__ __forceinline__ float es_tap(float x, float bw2) {
  float e;
  asm("{\n\t.reg .f32 t, s, nb;\n\t"
      "neg.f32 t, %1;\n\t"
      "fma.rn.f32 t, t, %1, 0f3F800000;\n\t"
      "sqrt.approx.ftz.f32 s, t;\n\t"
      "neg.f32 nb, %2;\n\t"
      "fma.rn.f32 s, %2, s, nb;\n\t"
      "ex2.approx.ftz.f32 s, s;\n\t"
      "max.f32 %0, s, 0f00000000;\n\t}"
      : "=f"(e) : "f"(x), "f"(bw2));
  return e;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ u64 dup2(float x) { return pack2(x, x); }

template <int OFF>
__device__ __forceinline__ u64 lds_u1(uint32_t addr) {
  u64 v;
  asm volatile("ld.shared.b64 %0, [%1+%2];" : "=l"(v) : "r"(addr), "n"(OFF));
  return v;
}

// Operands of one visibility for lane (a, pq): the row A[a][0..8] and the planes 2pq, 2pq+1 of B.
// Wavefront cost measured with tools/experiments/lds_patterns.cu: a 128-bit load is one wavefront per
// half-warp only when the lanes (2i, 2i+1) read the same 16 bytes (else one per quarter-warp), a 64-bit
// load is one wavefront per warp when its distinct words are conflict-free.  With lane = 4a + pq the two
// A loads cost 2 wavefronts each and the two B loads 1 each: 6 per visibility.
// Where a register window is folded to (see WarpAcc::fold).
struct FoldSink {
  float* grid;                                             // global plane stack
  int nu, nv, su0, sv0, P0, nplanes, plane_base, nslots;   // su0, sv0 already reduced to [0, nu), [0, nv)
};
__device__ __forceinline__ void red_add_v2(float* addr, u64 v) {
  const float2 f = unpack2(v);
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(f.x), "f"(f.y) : "memory");
}

// Row layouts of the operand stage, chosen so that BOTH consumers read them without bank conflicts:
//   A row (272 B): the four v residues 0..3 of u-row r at 16-byte unit (r >> 1) + 4 (r & 1), residues 4..7
//     eight units further.  The FFMA path reads one u-row per lane group a (lanes 2i, 2i+1 share the unit:
//     2 wavefronts per 128-bit load); the tensor path reads the same unit of four DIFFERENT visibilities
//     (row stride 17 units) for two neighbouring a in a quarter-warp: units pq + {0, 4} mod 8, conflict-free.
//   B row (80 B = 20 words): value (plane, re|im) = column n of the 8 x 16 operand, n = 8 (plane & 1) +
//     2 (plane >> 1) + (re|im).  Even planes at words {0..3, 16..19}, odd planes at words 4..11: the word
//     sets read from the rows of the four visibilities pq (row stride 20 words) are disjoint mod 32.
__device__ __forceinline__ constexpr int arow_unit(int r) { return (r >> 1) + 4 * (r & 1); }

// Tensor-core update of the register window by 8 visibilities that share it: C[64 x 16] += A[64 x 8] B[8 x 16]
// as 4 x 2 tiles of mma.m16n8k8 (legacy tensor path, 8 cycles per instruction and sub-partition measured with
// tools/experiments/mma_rate.cu -- 24 cycles per visibility with the split below, against 32 for the 16
// half-rate FFMA2).  TF32 operands carry 10 mantissa bits, so every operand is split x = hi + lo with hi the
// upper 19 bits and the three products lo*hi + hi*lo + hi*hi are accumulated from zero inside the tensor core,
// smallest first; the tile is then added to the window with IEEE float32 additions (the tensor core's own
// accumulator truncates).  Error per product ~ 2^-20.  Row m = 16 mt + g + 8 s of A is (u = g, v = 2 mt + s), so
// lane 4 g + t holds exactly the accumulators of the FFMA path (a = g, pq = t) and the folds are shared.
__device__ __forceinline__ void tf32_split(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x);  // the tensor core reads the upper 19 bits only
  lo = __float_as_uint(x - __uint_as_float(hi & 0xffffe000u));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// The first product of a chain: C = 0 comes from the zero register, so the accumulators need no clearing moves.
__device__ __forceinline__ void mma_tf32_first(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
               : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(0.0f));
}
// Two-term bfloat16 split (MMA = 2): x = hi + lo, both rounded to nearest, |x - hi - lo| <= 2^-18 |x|.  One m16n8k16 holds both terms of the 8 visibilities of a group
// along k -- slots 2t, 2t+1 = hi terms of the visibilities t and t + 4, slots 2t+8, 2t+9 = their lo terms -- so
// (A_hi + A_lo) B_hi is ONE instruction (b0 = b1) and (A_hi + A_lo) B_lo a second: 16 instead of 24 tensor
// instructions per group at the same issue rate (tools/experiments/mma_rate.cu).
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%8}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b));
}
__device__ __forceinline__ void mma_bf16_first(float (&d)[4], const uint32_t (&a)[4], uint32_t b) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%8}, {%9,%9,%9,%9};"
      : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b), "f"(0.0f));
}
__device__ __forceinline__ uint32_t bf16_rn_pair(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
template <int OFF>
__device__ __forceinline__ float2 lds_f2(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2+%3];" : "=f"(v.x), "=f"(v.y) : "r"(addr), "n"(OFF));
  return v;
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
template <int OFF>
__device__ __forceinline__ float lds_f32o(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(OFF));
  return v;
}

struct WarpOps {
  float4 a0, a1;
  u64 b0, b1;
};
template <int U, bool FACT>
__device__ __forceinline__ WarpOps warp_load(uint32_t pa, uint32_t pkv, uint32_t pb, uint32_t pb1) {
  WarpOps o;
  if (FACT) {  // pa: this lane's ku word, pkv: kv[0] of the same row
    const float ku = lds_f32(pa + U * kKRow);
    o.a0 = lds_f4<U * kKRow>(pkv);
    o.a1 = lds_f4<U * kKRow + 16>(pkv);
    o.a0.x *= ku; o.a0.y *= ku; o.a0.z *= ku; o.a0.w *= ku;
    o.a1.x *= ku; o.a1.y *= ku; o.a1.z *= ku; o.a1.w *= ku;
  } else {
    o.a0 = lds_f4<U * kARow>(pa);        // A row: v residues 0..3 of u-row a at 16 arow_unit(a), 4..7 at + 128
    o.a1 = lds_f4<U * kARow + 128>(pa);
  }
  o.b0 = lds_u1<U * kBRow>(pb);        // plane 2pq     (pb = row + brow_even(pq))
  o.b1 = lds_u1<U * kBRow>(pb1);       // plane 2pq + 1 (pb1 = row + 16 + 8 pq)
  return o;
}

struct WarpAcc {
  u64 c[8][2];  // [v residue][plane 2pq + h] = (re, im)
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int t = 0; t < 8; ++t) c[t][0] = c[t][1] = 0ull;
  }
  __device__ __forceinline__ void update(const WarpOps& o) {
    const float av[8] = {o.a0.x, o.a0.y, o.a0.z, o.a0.w, o.a1.x, o.a1.y, o.a1.z, o.a1.w};
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const u64 ad = dup2(av[t]);  // register-broadcast operand (R.F32), no instruction
      ffma2(c[t][0], ad, o.b0);
      ffma2(c[t][1], ad, o.b1);
    }
  }
  // pam: A unit of this lane's u-row in the row of visibility pq of the group; pbm0 / pbm1: this lane's column
  // (n = a) of the even / odd plane block in the same row.  Visibility pq + 4 sits four rows further.
  // Products of this lane's fragment rows for the visibilities pq and pq + 4 of a group: pam = factor row of
  // visibility pq, kuoff = byte offset of ku[a] in it.  The products are born in the registers the A fragments
  // need (aligned quads), so nothing is moved.
  // (v = 2m, 2m+1) of the visibilities pq (pp) and pq + 4 (qq) as packed float32 pairs
  // Addresses of a group (tensor path): pA = factor row of visibility pq, pK = pA + the byte offset of ku[a] in it --
  // the even-plane column n = a of the B row sits at the same offset of ITS row, i.e. at pK + kBStage -- and pB1 =
  // column n = a of the odd-plane block.  Three registers, advanced per group; every load is [register + immediate].
  static constexpr int kBStage = kBatch * kKRow;  // B stage behind the factor stage
  __device__ __forceinline__ static void load_product_pairs(uint32_t pam, uint32_t pK, u64 (&pp)[4], u64 (&qq)[4]) {
    const u64 kup = dup2(lds_f32o<0>(pK)), kuq = dup2(lds_f32o<4 * kKRow>(pK));
    const float4 p0 = lds_f4<16>(pam), p1 = lds_f4<32>(pam);
    const float4 q0 = lds_f4<4 * kKRow + 16>(pam), q1 = lds_f4<4 * kKRow + 32>(pam);
    pp[0] = fmul2(kup, pack2(p0.x, p0.y)); pp[1] = fmul2(kup, pack2(p0.z, p0.w));
    pp[2] = fmul2(kup, pack2(p1.x, p1.y)); pp[3] = fmul2(kup, pack2(p1.z, p1.w));
    qq[0] = fmul2(kuq, pack2(q0.x, q0.y)); qq[1] = fmul2(kuq, pack2(q0.z, q0.w));
    qq[2] = fmul2(kuq, pack2(q1.x, q1.y)); qq[3] = fmul2(kuq, pack2(q1.z, q1.w));
  }
  // Products, operand split and the float32 addition of the tiles work on packed pairs (FMUL2 / FADD2): the pairs
  // (v = 2 mt, 2 mt + 1) and (re, im) are exactly the register pairs of the fragments.
  __device__ __forceinline__ void update_mma(uint32_t pam, uint32_t pK, uint32_t pbm1) {
    u64 pp[4], qq[4];
    load_product_pairs(pam, pK, pp, qq);
    uint32_t bh[2][2], bl[2][2];
    tf32_split(lds_f32o<kBStage>(pK), bh[0][0], bl[0][0]);
    tf32_split(lds_f32o<kBStage + 4 * kBRow>(pK), bh[0][1], bl[0][1]);
    tf32_split(lds_f32o<0>(pbm1), bh[1][0], bl[1][0]);
    tf32_split(lds_f32o<4 * kBRow>(pbm1), bh[1][1], bl[1][1]);
    // two row tiles at a time: four independent accumulation chains (a dependent mma.sync issued after only
    // one other instruction of the tensor pipe waits for its accumulator)
#pragma unroll
    for (int mp = 0; mp < 4; mp += 2) {
      uint32_t ah[2][4], al[2][4];
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const u64 x = pp[mp + m], y = qq[mp + m];
        const u64 lp = fsub2(x, x & 0xffffe000ffffe000ull), lq = fsub2(y, y & 0xffffe000ffffe000ull);
        unpack2u(x, ah[m][0], ah[m][1]);  // hi = the raw bits: the tensor core reads the upper 19 only
        unpack2u(y, ah[m][2], ah[m][3]);
        unpack2u(lp, al[m][0], al[m][1]);
        unpack2u(lq, al[m][2], al[m][3]);
      }
      float d[2][2][4];
#pragma unroll
      for (int term = 0; term < 3; ++term)
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            if (term == 0) mma_tf32_first(d[m][nt], al[m], bh[nt][0], bh[nt][1]);
            else if (term == 1) mma_tf32(d[m][nt], ah[m], bl[nt][0], bl[nt][1]);
            else mma_tf32(d[m][nt], ah[m], bh[nt][0], bh[nt][1]);
          }
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          c[2 * (mp + m)][nt] = fadd2(c[2 * (mp + m)][nt], pack2(d[m][nt][0], d[m][nt][1]));
          c[2 * (mp + m) + 1][nt] = fadd2(c[2 * (mp + m) + 1][nt], pack2(d[m][nt][2], d[m][nt][3]));
        }
    }
  }
  // The same update with two-term bfloat16 operands (see mma_bf16): 16 tensor instructions per group.
  // x = (x0, x1), y = (y0, y1): h0 = (x0, y0), h1 = (x1, y1) as bf16 pairs rounded to nearest, l0 / l1 the bf16 pairs
  // of their residuals (a truncated hi term -- one PRMT, a mask for the residual -- measured 0.1 ms faster and twice
  // the error: 5.5e-6 against 2.8e-6 relative rms, tests/accuracy_modes.py).
  __device__ __forceinline__ static void bf16_split(u64 x, u64 y, uint32_t& h0, uint32_t& h1, uint32_t& l0, uint32_t& l1) {
    uint32_t x0, x1, y0, y1;
    unpack2u(x, x0, x1);
    unpack2u(y, y0, y1);
    h0 = bf16_rn_pair(__uint_as_float(x0), __uint_as_float(y0));
    h1 = bf16_rn_pair(__uint_as_float(x1), __uint_as_float(y1));
    const u64 lp = fsub2(x, pack2(__uint_as_float(h0 << 16), __uint_as_float(h1 << 16)));
    const u64 lq = fsub2(y, pack2(__uint_as_float(h0 & 0xffff0000u), __uint_as_float(h1 & 0xffff0000u)));
    const float2 rp = unpack2(lp), rq = unpack2(lq);
    l0 = bf16_rn_pair(rp.x, rq.x);
    l1 = bf16_rn_pair(rp.y, rq.y);
  }
  __device__ __forceinline__ void update_mma_bf16(uint32_t pam, uint32_t pK, uint32_t pbm1) {
    u64 pp[4], qq[4];
    load_product_pairs(pam, pK, pp, qq);
    uint32_t bh[2], bl[2];
    // B: (even block of visibility pq, odd block of pq) and the same of visibility pq + 4
    bf16_split(pack2(lds_f32o<kBStage>(pK), lds_f32o<0>(pbm1)),
                   pack2(lds_f32o<kBStage + 4 * kBRow>(pK), lds_f32o<4 * kBRow>(pbm1)), bh[0], bh[1], bl[0], bl[1]);
#pragma unroll
    for (int mp = 0; mp < 4; mp += 2) {
      uint32_t af[2][4];
#pragma unroll
      for (int m = 0; m < 2; ++m) bf16_split(pp[mp + m], qq[mp + m], af[m][0], af[m][1], af[m][2], af[m][3]);
      float d[2][2][4];
#pragma unroll
      for (int term = 0; term < 2; ++term)
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            if (term == 0) mma_bf16_first(d[m][nt], af[m], bl[nt]);
            else mma_bf16(d[m][nt], af[m], bh[nt]);
          }
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          c[2 * (mp + m)][nt] = fadd2(c[2 * (mp + m)][nt], pack2(d[m][nt][0], d[m][nt][1]));
          c[2 * (mp + m) + 1][nt] = fadd2(c[2 * (mp + m) + 1][nt], pack2(d[m][nt][2], d[m][nt][3]));
        }
    }
  }
  // the lane owns row i of the window (all 8 v residues) for the plane pair pq
  template <int T0>
  __device__ __forceinline__ void fold_pairs(const FoldSink& k, float* row0, float* row1, bool ok0, bool ok1, int wj) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      constexpr int dummy = 0;
      (void)dummy;
      const int ta = (T0 + 2 * q) & 7, tb = (T0 + 2 * q + 1) & 7;
      const int ja = wj + ((ta - wj) & 7), jb = wj + ((tb - wj) & 7);
      int ga = k.sv0 + ja, gb = k.sv0 + jb;
      ga -= ga >= k.nv ? k.nv : 0;
      gb -= gb >= k.nv ? k.nv : 0;
      const bool adj = gb == ga + 1;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        if (!(h ? ok1 : ok0)) continue;
        float* row = h ? row1 : row0;
        const u64 va = c[ta][h], vb = c[tb][h];
        if ((va | vb) == 0ull) continue;
        if (adj) {
          const float2 x = unpack2(va), y = unpack2(vb);
          red_add_v4(row + 2 * ga, x.x, x.y, y.x, y.y);
        } else {
          if (va != 0ull) red_add_v2(row + 2 * ga, va);
          if (vb != 0ull) red_add_v2(row + 2 * gb, vb);
        }
      }
    }
  }
  // The window goes straight to the global plane stack (no sub-grid in shared memory: the records are sorted by
  // window, so a warp leaves a window about once per lattice position and item; measured 10.8 -> 10.0 ms in round 1
  // against a shared sub-grid, which halves the resident warps).
  __device__ __forceinline__ void fold(const FoldSink& k, int i, int wj, int pq) {
    int gi = k.su0 + i;
    gi -= gi >= k.nu ? k.nu : 0;
    const int plane = k.P0 + 2 * pq, slot = plane - k.plane_base;
    const bool ok0 = plane < k.nplanes && slot >= 0 && slot < k.nslots;
    const bool ok1 = plane + 1 < k.nplanes && slot + 1 >= 0 && slot + 1 < k.nslots;
    float* row0 = k.grid + 2 * (((int64_t)slot * k.nu + gi) * k.nv);
    float* row1 = row0 + 2 * (int64_t)k.nu * k.nv;
    // Residue t sits at window column (t - wj) & 7, i.e. global column parity = parity(sv0 + t): the
    // residue pairs (t, t+1) that start on an even global column are 16-byte aligned neighbours unless the
    // window wraps between them, so most of the window goes out as 128-bit reductions (one sector per
    // request lane instead of two 8-byte requests to the same sector).
    if (k.sv0 & 1) fold_pairs<1>(k, row0, row1, ok0, ok1, wj);
    else fold_pairs<0>(k, row0, row1, ok0, ok1, wj);
    clear();
  }
};

struct WarpCells {
  int i, wj;
  __device__ __forceinline__ void set(uint32_t org, int a) {
    const int wi = org & 31;
    wj = (org >> 5) & 31;
    i = wi + ((a - wi) & 7);
  }
};

template <int U, bool CHECK>
__device__ __forceinline__ void warp_consume1(uint32_t pa, uint32_t pb, uint32_t pb1, uint32_t porg, uint32_t moved8, const WarpOps& cur,
                                              WarpOps& nxt, WarpAcc& acc, const FoldSink& sg, WarpCells& cells, int a, int pq) {
  nxt = warp_load<U + 1, false>(pa, 0u, pb, pb1);
  if (CHECK && ((moved8 >> U) & 1u)) {
    acc.fold(sg, cells.i, cells.wj, pq);
    cells.set(lds_u32(porg + 4 * U), a);
  }
  acc.update(cur);
}
template <bool CHECK>
__device__ __forceinline__ void warp_consume8(uint32_t pa, uint32_t pb, uint32_t pb1, uint32_t porg, uint32_t moved8, WarpOps& x, WarpAcc& acc,
                                              const FoldSink& sg, WarpCells& cells, int a, int pq) {
  WarpOps y;
  warp_consume1<0, CHECK>(pa, pb, pb1, porg, moved8, x, y, acc, sg, cells, a, pq);
  warp_consume1<1, CHECK>(pa, pb, pb1, porg, moved8, y, x, acc, sg, cells, a, pq);
  warp_consume1<2, CHECK>(pa, pb, pb1, porg, moved8, x, y, acc, sg, cells, a, pq);
  warp_consume1<3, CHECK>(pa, pb, pb1, porg, moved8, y, x, acc, sg, cells, a, pq);
  warp_consume1<4, CHECK>(pa, pb, pb1, porg, moved8, x, y, acc, sg, cells, a, pq);
  warp_consume1<5, CHECK>(pa, pb, pb1, porg, moved8, y, x, acc, sg, cells, a, pq);
  warp_consume1<6, CHECK>(pa, pb, pb1, porg, moved8, x, y, acc, sg, cells, a, pq);
  warp_consume1<7, CHECK>(pa, pb, pb1, porg, moved8, y, x, acc, sg, cells, a, pq);
}

// One warp = one work item.  MMA = 1: tensor-core tiles for the groups that stay inside one window; 0: FFMA2
// rank-1 updates throughout (A/B reference).
template <int MMA>
__device__ __forceinline__ void grid_warp_item(const Item& item, const RecP* __restrict__ recs, const GridGeom& g,
                                               float* __restrict__ grid, int plane_base, int nslots) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31;
  constexpr bool FACT = MMA != 0;
  constexpr int ROW = FACT ? kKRow : kARow;
  const uint32_t sA = smem_u32(smem_raw);
  const uint32_t sB = sA + kBatch * ROW;
  const uint32_t sOrg = sB + kBatch * kBRow;

  const int nrec = item.nrec;
  const int b_begin = 0, b_end = (nrec + kBatch - 1) / kBatch;
  const int64_t base = item.rec_begin;

  bool cur_valid = lane < nrec;
  RecP cur = load_recp(recs, base + lane, cur_valid);
  const int P0 = item.win * g.win_step;
  int su0 = item.tu * kTile - kHalo + g.nu / 2;  // shifted -> FFT order: (s + nu/2) mod nu
  int sv0 = item.tv * kTile - kHalo + g.nv / 2;
  su0 -= su0 >= g.nu ? g.nu : 0;  // reduced once here; su0 + i < 2 nu
  sv0 -= sv0 >= g.nv ? g.nv : 0;
  const FoldSink sink{grid, g.nu, g.nv, su0, sv0, P0, g.nplanes, plane_base, nslots};

  // producer constants
  const int W = g.W;
  const float bw2 = g.beta_w * 1.4426950408889634f;
  const float two_over_w = 2.0f / (float)W;
  const int snap = 9 - W;  // windows of 8 cells hold the supports of `snap` consecutive origins
  const uint32_t snap_rcp = (256u + snap - 1) / snap;
  const bool wstack = g.do_wstacking != 0;
  uint32_t carry = 0xffffffffu;
  const uint32_t myA = sA + lane * ROW, myB = sB + lane * kBRow;
  // consumer constants
  const int a = lane >> 2, pq = lane & 3;
  const uint32_t kuoff = 4 * (a < 4 ? a : 12 + a);
  const uint32_t pa0 = sA + (FACT ? kuoff : 16 * arow_unit(a)), pkv0 = sA + 16;
  const uint32_t pb0 = sB + (pq < 2 ? pq * 8 : 48 + pq * 8), pb10 = sB + 16 + pq * 8;
  // tensor path: rows of visibilities pq and pq + 4 of a group, column n = a of the B blocks
  const uint32_t pam0 = sA + pq * kKRow;
  const uint32_t pbm10 = sB + pq * kBRow + 4 * (4 + a);  // (the even block's column sits at pam0 + kuoff + kBStage)
  WarpAcc acc;
  acc.clear();
  WarpCells cells;
  cells.set(0u, a);
  __syncwarp();

  for (int b = b_begin; b < b_end; ++b) {
    const int kn = (b + 1) * kBatch + lane;
    const bool nxt_valid = kn < nrec;
    // (This register prefetch is spilled at 80 registers and the spill store waits for the load -- 9 % of the stall
    // samples sit there -- yet both alternatives measured slower on configs[1]: cp.async into a shared-memory record
    // buffer 7.33 ms, prefetch.global.L1 + a load at the end of the batch 7.44 ms, against 6.91 ms.)
    const RecP nxt = load_recp(recs, base + kn, nxt_valid);
    // ---------------- produce: lane = visibility ----------------
    uint32_t moved;
    {
      // unpack: origin inside the tile, first plane, 0.18 fixed-point fractions; the register window's origin
      // is the support origin snapped to the lattice (i / snap by multiplication, exact for i < 8)
      const uint32_t glo = (uint32_t)cur.geo, ghi = (uint32_t)(cur.geo >> 32);
      const int i0 = glo & 7, j0 = (glo >> 3) & 7, p0 = (glo >> 6) & 15;
      const float cur_fu = (float)((glo >> 10) & 0x3ffffu) * (1.0f / 262144.0f);
      const float cur_fv = (float)((uint32_t)(cur.geo >> 28) & 0x3ffffu) * (1.0f / 262144.0f);
      const float cur_fw = (float)(ghi >> 14) * (1.0f / 262144.0f);
      const uint32_t wi = ((i0 * snap_rcp) >> 8) * snap, wj = ((j0 * snap_rcp) >> 8) * snap;
      const uint32_t cxy = window_origin((wi | (wj << 5)) << 16, cur_valid, lane, carry, snap, &moved);
      // Taps by residue class r: the tap of cell r sits da = (r - i0) mod 8 cells from the first one, at
      // x = (f + da) 2/W - 1; taps outside the support have |x| >= 1 and are masked by the sign of 1 - x^2
      // (branch-free; at |x| = 1 exactly the kernel is exp(-beta W) ~ 1e-7 of its peak, i.e. nothing).
      const float i0f = (float)i0, j0f = (float)j0, p0f = (float)p0;
      const float xu0 = fmaf(cur_fu - i0f, two_over_w, -1.0f), xv0 = fmaf(cur_fv - j0f, two_over_w, -1.0f);
      const float xw0 = fmaf(cur_fw - p0f, two_over_w, -1.0f);
      float ku[8], kv[8], kw[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float c0 = (float)r * two_over_w, c1 = (float)(r + 8) * two_over_w;
        ku[r] = es_tap(xu0 + (r < i0 ? c1 : c0), bw2);
        kv[r] = es_tap(xv0 + (r < j0 ? c1 : c0), bw2);
        kw[r] = wstack ? es_tap(xw0 + c0, bw2) : (r == 0 ? 1.0f : 0.0f);
      }
      if (FACT) {
        sts_f4(myA, ku[0], ku[1], ku[2], ku[3]);
        sts_f4(myA + 16, kv[0], kv[1], kv[2], kv[3]);
        sts_f4(myA + 32, kv[4], kv[5], kv[6], kv[7]);
        sts_f4(myA + 64, ku[4], ku[5], ku[6], ku[7]);
      } else {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          sts_f4(myA + 16 * arow_unit(r), ku[r] * kv[0], ku[r] * kv[1], ku[r] * kv[2], ku[r] * kv[3]);
          sts_f4(myA + 16 * arow_unit(r) + 128, ku[r] * kv[4], ku[r] * kv[5], ku[r] * kv[6], ku[r] * kv[7]);
        }
      }
      const u64 v2 = cur_valid ? pack2(cur.re, cur.im) : 0ull;
#pragma unroll
      for (int q = 0; q < 8; q += 4) {
        sts_u2(myB + q * 16, fmul2(v2, dup2(kw[q])), fmul2(v2, dup2(kw[q + 2])));               // even planes
        sts_u2(myB + 16 + q * 4, fmul2(v2, dup2(kw[q + 1])), fmul2(v2, dup2(kw[q + 3])));  // odd planes
      }
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(sOrg + 4 * lane), "r"(cxy) : "memory");
    }
    __syncwarp();
    // ---------------- consume: lane = (a, bh, ph) ----------------
    if constexpr (MMA != 0) {
      // Tensor path.  Every key is padded to whole k-groups (bin.cu), so a window moves at a group start: fold there
      // and keep the group on the tensor cores.  The three group addresses live in registers the compiler may not
      // re-derive (it used to rebuild them from the thread index for every group: 20 of ~110 instructions).
      uint32_t pA = pam0, pK = pam0 + kuoff, pB1 = pbm10;
#pragma unroll 1
      for (int k0 = 0; k0 < kBatch; k0 += 8) {
        asm volatile("" : "+r"(pA), "+r"(pK), "+r"(pB1));
        const uint32_t m8 = (moved >> k0) & 0xffu;
        if (m8 & 1u) {
          acc.fold(sink, cells.i, cells.wj, pq);
          cells.set(lds_u32(sOrg + 4 * k0), a);
        }
        if ((m8 & ~1u) == 0) {
          if (MMA >= 2) acc.update_mma_bf16(pA, pK, pB1);
          else acc.update_mma(pA, pK, pB1);
        } else {
          // a window moves inside the group (records that do not come from bin.cu's padded keys): rank-1 updates
          const uint32_t pa = pa0 + k0 * ROW, pkv = pkv0 + k0 * ROW, pb = pb0 + k0 * kBRow, pb1 = pb10 + k0 * kBRow;
          WarpOps x = warp_load<0, FACT>(pa, pkv, pb, pb1);
#pragma unroll 1
          for (int u = 0; u < 8; ++u) {
            const WarpOps nxt = warp_load<1, FACT>(pa + u * ROW, pkv + u * ROW, pb + u * kBRow, pb1 + u * kBRow);
            if (u && ((m8 >> u) & 1u)) {
              acc.fold(sink, cells.i, cells.wj, pq);
              cells.set(lds_u32(sOrg + 4 * (k0 + u)), a);
            }
            acc.update(x);
            x = nxt;
          }
        }
        pA += 8 * kKRow;
        pK += 8 * kKRow;
        pB1 += 8 * kBRow;
      }
    } else {
      WarpOps x = warp_load<0, false>(pa0, 0u, pb0, pb10);
#pragma unroll 1
      for (int k0 = 0; k0 < kBatch; k0 += 8) {
        const uint32_t pa = pa0 + k0 * ROW, pb = pb0 + k0 * kBRow, pb1 = pb10 + k0 * kBRow;
        const uint32_t porg = sOrg + 4 * k0;
        const uint32_t m8 = (moved >> k0) & 0xffu;
        // most groups stay inside one register window: branch-free fast path.  The operands of visibility
        // k0 + 8 are prefetched at the end of a group; past the last group that reads the rows after the
        // stage (still inside the CTA's shared memory: kSlack), which are discarded.
        if (m8 == 0) {
          warp_consume8<false>(pa, pb, pb1, porg, m8, x, acc, sink, cells, a, pq);
        } else {
          // a window moves inside this group: rolled loop, so that the fold is not inlined eight times
          // (the unrolled variant doubled the code size and a third of the stalls were instruction fetches)
#pragma unroll 1
          for (int u = 0; u < 8; ++u) {
            const WarpOps nxt = warp_load<1, false>(pa + u * ROW, 0u, pb + u * kBRow, pb1 + u * kBRow);
            if ((m8 >> u) & 1u) {
              acc.fold(sink, cells.i, cells.wj, pq);
              cells.set(lds_u32(porg + 4 * u), a);
            }
            acc.update(x);
            x = nxt;
          }
        }
      }
    }
    __syncwarp();
    cur = nxt;
    cur_valid = nxt_valid;
  }
  acc.fold(sink, cells.i, cells.wj, pq);
}

// The work items of the pass (bin.cu): the list of their first jobs; an item runs up to the next job that starts one.
struct ItemList {
  const Job* jobs;
  const uint32_t* items;
  uint32_t* counters;  // [0] number of items, [1] next item to hand out
  const BinHeader* hdr;
  int win_first, win_last;
};

// Resident one-warp CTAs per SM the register allocation aims at: 24 (80 registers) for the tensor paths, measured
// against 20 / 28 / 32 (3xTF32 7.56 / 7.33 / 7.47 ms, bf16 7.05 / 6.92 ms at 20 / 24 on configs[1]); 20 for FFMA2.
constexpr int warp_minb(int mma) { return mma ? 24 : 20; }
template <int MMA>
__global__ void __launch_bounds__(32, warp_minb(MMA)) grid_warp_kernel(ItemList il, const RecP* __restrict__ recs, GridGeom g,
                                                            float* __restrict__ grid, int plane_base, int nslots) {
  if (il.hdr->overflow) return;
  const uint32_t n = il.counters[0], njobs = il.hdr->njobs;
  const int lane = threadIdx.x & 31;
  for (;;) {
    // items are handed out dynamically: their sizes differ by orders of magnitude
    uint32_t i = 0;
    if (lane == 0) i = atomicAdd(il.counters + 1, 1u);
    i = __shfl_sync(0xffffffffu, i, 0);
    if (i >= n) break;
    const uint32_t ja = il.items[i];
    const Job a = il.jobs[ja];
    if ((int)a.win < il.win_first || (int)a.win > il.win_last) continue;
    // the item ends before the next job that starts one
    uint32_t jb = ja + 1;
    for (;; jb += 32) {
      const uint32_t k = jb + lane;
      const bool stop = k >= njobs || (il.jobs[k].ngroups & kJobStartsItem);
      const unsigned m = __ballot_sync(0xffffffffu, stop);
      if (m) {
        jb += __ffs(m) - 1;
        break;
      }
    }
    const Job b = il.jobs[jb - 1];
    Item item;
    item.rec_begin = (int64_t)a.group_begin * kGroup;
    item.nrec = (int32_t)((b.group_begin + (b.ngroups & ~kJobStartsItem) - a.group_begin) * kGroup);
    item.tu = a.tu;
    item.tv = a.tv;
    item.win = a.win;
    grid_warp_item<MMA>(item, recs, g, grid, plane_base, nslots);
    __syncwarp();
  }
}

// Wide supports (W > 8, only reached for very small epsilon): plain scatter with global atomics.
template <typename R>
__global__ void __launch_bounds__(128) grid_wide_kernel(JobList jl, const Rec<R>* __restrict__ recs, GridGeom g, R* __restrict__ grid,
                                                        int plane_base, int nslots) {
  uint32_t jb, je;
  job_range(jl, g, &jb, &je);
  for (uint32_t jix = jb + blockIdx.x; jix < je; jix += gridDim.x) {
  const Item item = job_item(jl.jobs[jix], g);
  const int W = g.W;
  const int su0 = item.tu * kTile - kHalo + g.nu / 2, sv0 = item.tv * kTile - kHalo + g.nv / 2;
  for (int k = threadIdx.x; k < item.nrec; k += blockDim.x) {
    const Rec<R> rec = load_rec(recs, item.rec_begin + k, true);
    if (rec.re == 0 && rec.im == 0) continue;
    const int i0 = rec.cell & 31, j0 = (rec.cell >> 5) & 31, p0 = rec.cell >> 10;
    // kernel weights in the precision of the grids: float32 evaluation of a wide ES kernel loses ~1e-5
    R fu, fv, fw, bw;
    if constexpr (sizeof(R) == 4) {
      fu = rec.fu; fv = rec.fv; fw = rec.fw; bw = g.beta_w;
    } else {
      fu = from_fix32(rec.fu); fv = from_fix32(rec.fv); fw = from_fix32(rec.fw); bw = g.beta_wd;
    }
    R kv[kMaxSupport];
    for (int b = 0; b < W; ++b) kv[b] = es_kernel(((R)2 * (fv + b) - W) / W, bw);
    for (int c = 0; c < (g.do_wstacking ? W : 1); ++c) {
      int plane = p0 + c, slot = plane - plane_base;
      if (plane >= g.nplanes || slot < 0 || slot >= nslots) continue;
      R kw = g.do_wstacking ? es_kernel(((R)2 * (fw + c) - W) / W, bw) : (R)1;
      for (int a = 0; a < W; ++a) {
        R ku = es_kernel(((R)2 * (fu + a) - W) / W, bw) * kw;
        int gi = (su0 + i0 + a) % g.nu;
        if (gi < 0) gi += g.nu;
        R tr = rec.re * ku, ti = rec.im * ku;
        for (int b = 0; b < W; ++b) {
          int gj = (sv0 + j0 + b) % g.nv;
          if (gj < 0) gj += g.nv;
          R* dst = grid + 2 * (((int64_t)slot * g.nu + gi) * g.nv + gj);
          atomicAdd(dst, tr * kv[b]);
          atomicAdd(dst + 1, ti * kv[b]);
        }
      }
    }
  }
  }
}

}  // namespace

cip_status run_gridder_tc(cip_ctx* ctx, int slot, cudaStream_t stream, const GridGeom& g, int win_first, int win_last, float* grid,
                          int plane_base, int nslots);

namespace {

template <int MMA>
cip_status launch_warp(cip_ctx* ctx, cudaStream_t stream, const ItemList& jl, const RecP* recs, const GridGeom& g, float* grid,
                       int plane_base, int nslots) {
  auto kern = grid_warp_kernel<MMA>;
  constexpr int smem = warp_smem_bytes(MMA != 0);
  CIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
  kern<<<ctx->sm_count * warp_minb(MMA), 32, smem, stream>>>(jl, recs, g, grid, plane_base, nslots);
  return CIP_OK;
}

template <typename R>
cip_status launch(cip_ctx* ctx, int slot, cudaStream_t stream, const GridGeom& g, int win_first, int win_last, void* grid,
                  int plane_base, int nslots) {
  const BinWork& w = ctx->work[slot];
  const JobList jl{w.jobs.as<Job>(), w.win_job_start.as<uint32_t>(), w.hdr.as<BinHeader>(), win_first, win_last};
  const Rec<R>* recs = w.recs.as<Rec<R>>();
  if (g.W <= kFastSupport) {
    if constexpr (sizeof(R) == 4) {
      // float32: the register-window kernel (3xTF32 mma.sync tiles, float32 additions) on the work items of the
      // pass; CIP_GRID_MMA=0 selects its FFMA2 form (A/B).  CIP_GRID_TC=1 selects the tcgen05 kernel of grid_tc.cu
      // instead -- correct, and the tensor pipe itself is fast there, but feeding its A operand through tensor
      // memory costs more shared-memory bandwidth and issue overhead than the whole of this kernel.
      if (use_tc_gridder())
        return run_gridder_tc(ctx, slot, stream, g, win_first, win_last, static_cast<float*>(grid), plane_base, nslots);
      uint32_t* counters = w.items.as<uint32_t>() + w.cap_jobs;
      CIP_CHECK_CUDA(cudaMemsetAsync(counters + 1, 0, sizeof(uint32_t), stream));  // the same items may be gridded again (plane blocks)
      const ItemList il{w.jobs.as<Job>(), w.items.as<uint32_t>(), counters, w.hdr.as<BinHeader>(), win_first, win_last};
      // Operand form of the tensor path: two-term bfloat16 (2.8e-6 relative rms arithmetic error) where the plan's
      // kernel is coarser than that anyway, 3xTF32 (the float32 floor, 2e-6 against the FFMA2 form's 1.9e-6) otherwise.
      const int mma = getenv("CIP_GRID_MMA") ? atoi(getenv("CIP_GRID_MMA")) : (g.bf16_operands ? 2 : 1);
      const RecP* rp = reinterpret_cast<const RecP*>(recs);
      float* gp = static_cast<float*>(grid);
      if (mma == 0) CIP_TRY((launch_warp<0>(ctx, stream, il, rp, g, gp, plane_base, nslots)));
      else if (mma == 2) CIP_TRY((launch_warp<2>(ctx, stream, il, rp, g, gp, plane_base, nslots)));
      else CIP_TRY((launch_warp<1>(ctx, stream, il, rp, g, gp, plane_base, nslots)));
    } else {
      // float64: producer / consumer warp pair ("residue owner")
      constexpr size_t smem = grid_smem_bytes<R>();
      auto kern = grid_kernel<R>;
      CIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      CIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                          (int)cudaSharedmemCarveoutMaxShared));
      kern<<<ctx->sm_count * 8, kGridThreads, smem, stream>>>(jl, recs, g, static_cast<R*>(grid), plane_base, nslots);
    }
  } else {
    grid_wide_kernel<R><<<ctx->sm_count * 16, 128, 0, stream>>>(jl, recs, g, static_cast<R*>(grid), plane_base, nslots);
  }
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

}  // namespace

bool use_tc_gridder() {
  const char* e = getenv("CIP_GRID_TC");
  return e && atoi(e) != 0;
}

cip_status run_gridder(cip_ctx* ctx, int slot, cudaStream_t stream, const GridGeom& g, int fp64, int win_first, int win_last,
                       void* grid, int plane_base, int nslots) {
  NvtxRange nvtx("cip:gridding");
  if (fp64) return launch<double>(ctx, slot, stream, g, win_first, win_last, grid, plane_base, nslots);
  return launch<float>(ctx, slot, stream, g, win_first, win_last, grid, plane_base, nslots);
}

}  // namespace cip
