// Kernel 2, float32, supports W <= 8: w-stacked gridding on the 5th-generation tensor cores (tcgen05, accumulators
// in tensor memory) -- stage G2 of SURVEY.md section 8a, the hot loop of ducc0.wgridder.ms2dirty (reference call
// site src/ska_sdp_cip/invert.py:170-183).
//
// The update of one 8 x 8 x 8 accumulator window (u, v, w plane) by a visibility is the rank-1 product a (x) b,
// a[(u, v)] = ku[u] kv[v] (64 values), b[(plane, re|im)] = V kw[plane] (16 values); the K visibilities of one job
// (= one window, see bin.cu) make it a GEMM  C[64 x 16] += A[64 x K] B[K x 16].  float32 accuracy on TF32 tensor
// cores (which truncate operands to 11 significant bits -- tools/experiments/tcgen05_probe.cu) takes split
// operands x = hi + lo and the products hi hi + lo hi + hi lo (+ lo lo); they are ONE tcgen05.mma per 8
// visibilities:  D[128 x 32] += [Ah; Al][128 x 8] [Bh | Bl][8 x 32], kind::tf32, M = 128, N = 32, K = 8, 16 cycles
// (measured), with A read from tensor memory -- shared-memory A operands cost 23 .. 40 cycles per mma because the
// 64 products per visibility would cross the shared-memory port twice.
//
// One persistent CTA per SM, warp-specialised:
//   producers  lane = visibility: 16-byte record, the separable ES kernel (8 u-, 8 v-, 8 w-taps by residue class of
//              the window, branch-free), the factors ku, kv to shared memory and the split 32-column B operand in the
//              K-major core-matrix layout the tensor core reads;
//   builders   thread = row of A = tensor-memory lane (rows 2c, 2c+1 = hi, lo of cell c): products ku kv of 32
//              visibilities, split, tcgen05.st;
//   issuer     one thread: one tcgen05.mma per k-group, tcgen05.commit to recycle operand stages and to hand
//              finished windows to the epilogue;
//   epilogue   tcgen05.ld of a finished window (8 accumulator windows rotate in tensor memory), hi + lo parts
//              summed, added to the global plane stack with red.global.add.v2.f32.
// Hand-over by mbarriers only; no __syncthreads after the prologue.
#include <cstdlib>
#include <vector>

#include "geom.cuh"

namespace cip {

namespace {

constexpr int kTcStages = 8;     // operand stages; one stage = one batch of 32 records = 4 k-groups
constexpr int kTcAccs = 8;       // accumulator windows in tensor memory (32 columns each)
constexpr int kTcProducers = 7;
constexpr int kTcJobCost = 16;   // per-job work in k-group equivalents, for the split of the job list over the CTAs
constexpr int kTcWarps = 4 + 4 + 1 + kTcProducers;
constexpr int kTcThreads = 32 * kTcWarps;
// B operand of one k-group (N = 32, K = 8, K-major, no swizzle): core matrix (n / 8, k / 4) = 8 rows of 16 bytes at
// (n / 8) * kBSbo + (k / 4) * kBLbo; the strides are padded so that the producers' 32-bit stores (lane = visibility:
// k / 4, k % 4 and the k-group vary over the warp) touch 32 different banks.
constexpr uint32_t kBLbo = 144, kBSbo = 288, kBGroup = 1184;
constexpr uint32_t kStageF = 2048;                                  // ku[8][32] and kv[8][32] floats
constexpr uint32_t kStageBytes = kStageF + 4 * kBGroup;             // 6784
constexpr uint32_t kTcSmem = kTcStages * kStageBytes + 1024;        // + barriers and bookkeeping

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}
// For the roles that usually wait long (producers for a free stage, the epilogue for a finished window): back off
// between polls so that the polling does not take issue slots from the warps that work.
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity, uint32_t ns) {
  uint32_t ok;
  for (;;) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) break;
    __nanosleep(ns);
  }
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc)
               : "memory");
}
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred)::"memory");
  return pred;
}
// kind::tf32, float32 accumulation, M = 128, N = 32, A and B K-major (bit layout: cute/arch/mma_sm100_desc.hpp)
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);
__device__ __forceinline__ uint64_t make_bdesc(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)(kBLbo >> 4) << 16) | ((uint64_t)(kBSbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,"
      "%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
      "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]),
      "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,"
      "%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void sts32(uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }
// ES tap exp(beta_w (sqrt(1 - x^2) - 1)), 0 where 1 - x^2 <= 0; pure PTX so that the mask stays a select
__device__ __forceinline__ float es_tap(float x, float bw2) {
  float e;
  asm("{\n\t.reg .f32 t, s, nb;\n\t.reg .pred p;\n\t"
      "neg.f32 t, %1;\n\t"
      "fma.rn.f32 t, t, %1, 0f3F800000;\n\t"
      "setp.gt.f32 p, t, 0f00000000;\n\t"
      "sqrt.approx.ftz.f32 s, t;\n\t"
      "neg.f32 nb, %2;\n\t"
      "fma.rn.f32 s, %2, s, nb;\n\t"
      "ex2.approx.ftz.f32 s, s;\n\t"
      "selp.f32 %0, s, 0f00000000, p;\n\t}"
      : "=f"(e) : "f"(x), "f"(bw2));
  return e;
}
__device__ __forceinline__ float tf32_lo(float x) { return x - __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

struct TcShared {
  unsigned long long f_full[kTcStages], a_full[kTcStages], s_free[kTcStages], acc_full[kTcAccs], acc_free[kTcAccs];
  uint32_t tmem_base;
  uint32_t j0, j1;
};

// PROF: per-role cycle counters (lane 0 of one warp per role) into prof[cta][16]: development aid, CIP_TC_PROF=1.
template <bool PROF>
__global__ void __launch_bounds__(kTcThreads, 1) grid_tc_kernel(const Job* __restrict__ jobs, const uint32_t* __restrict__ win_job_start,
                                                                const BinHeader* __restrict__ hdr, const RecP* __restrict__ recs,
                                                                GridGeom g, float* __restrict__ grid, int win_first, int win_last,
                                                                int plane_base, int nslots, int job_cost, uint32_t relax_ns, int dbg, long long* __restrict__ prof) {
  extern __shared__ __align__(1024) unsigned char smem[];
  long long t_wait0 = 0, t_wait1 = 0, t_begin = 0, t_alu = 0, t_mma = 0;
  auto timed_wait = [&](uint32_t bar, uint32_t parity, long long& acc, uint32_t relax_ns = 0) {
    const long long t = PROF ? clock64() : 0;
    if (relax_ns) mbar_wait_relaxed(bar, parity, relax_ns);
    else mbar_wait(bar, parity);
    if (PROF) acc += clock64() - t;
  };
  TcShared* sh = reinterpret_cast<TcShared*>(smem + kTcStages * kStageBytes);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (hdr->overflow) return;

  // ---- this CTA's share: a contiguous range of jobs holding ~ 1 / gridDim.x of the k-groups --------------------
  if (tid == 0) {
    const uint32_t jb = win_job_start[win_first - g.win_begin], je = win_job_start[win_last + 1 - g.win_begin];
    uint32_t j0 = jb, j1 = jb;
    if (je > jb) {
      // cost of the jobs before job m: their k-groups plus kTcJobCost group-equivalents of per-job work (accumulator
      // turn-around, epilogue) -- sparse regions of the uv plane hold many jobs of one or two groups
      const uint64_t gb = jobs[jb].group_begin, ge = (uint64_t)jobs[je - 1].group_begin + (jobs[je - 1].ngroups & ~kJobStartsItem);
      const uint64_t total = (ge - gb) + (uint64_t)job_cost * (je - jb);
      const uint64_t lo = total * blockIdx.x / gridDim.x, hi = total * (blockIdx.x + 1) / gridDim.x;
      auto first_at_or_after = [&](uint64_t x) {  // first job whose preceding cost is >= x
        uint32_t a = jb, b = je;
        while (a < b) {
          const uint32_t m = a + (b - a) / 2;
          if ((jobs[m].group_begin - gb) + (uint64_t)job_cost * (m - jb) < x) a = m + 1;
          else b = m;
        }
        return a;
      };
      j0 = first_at_or_after(lo);
      j1 = blockIdx.x + 1 == gridDim.x ? je : first_at_or_after(hi);
    }
    sh->j0 = j0;
    sh->j1 = j1;
    for (int s = 0; s < kTcStages; ++s) {
      mbar_init(smem_u32(&sh->f_full[s]), 1);
      mbar_init(smem_u32(&sh->a_full[s]), 4);
      mbar_init(smem_u32(&sh->s_free[s]), 1);
    }
    for (int a = 0; a < kTcAccs; ++a) {
      mbar_init(smem_u32(&sh->acc_full[a]), 1);
      mbar_init(smem_u32(&sh->acc_free[a]), 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint32_t j0 = sh->j0, j1 = sh->j1;
  if (j0 >= j1) return;  // uniform over the CTA
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sh->tmem_base)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tbase = sh->tmem_base;
  const uint32_t gbeg = jobs[j0].group_begin;
  const uint32_t gend = jobs[j1 - 1].group_begin + (jobs[j1 - 1].ngroups & ~kJobStartsItem);
  const uint32_t nbatch = (gend - gbeg + 3) / 4;
  const uint32_t sbase = smem_u32(smem);
  if (PROF) t_begin = clock64();

  if (warp >= 9) {
    // ================= producers: lane = visibility =================
    const int p = warp - 9;
    const int W = g.W;
    const float bw2 = g.beta_w * 1.4426950408889634f;
    const float two_over_w = 2.0f / (float)W;
    const bool wstack = g.do_wstacking != 0;
    const uint64_t rec_end = (uint64_t)gend * kGroup;
    uint64_t idx = (uint64_t)(gbeg + 4 * p) * kGroup + lane;
    float4 cur = idx < rec_end ? __ldg(reinterpret_cast<const float4*>(recs + idx)) : make_float4(0.f, 0.f, 0.f, 0.f);
    const uint32_t kgrp = lane >> 3, kh = (lane >> 2) & 1, k4 = lane & 3;
    for (uint32_t i = p; i < nbatch; i += kTcProducers) {
      const uint32_t s = i % kTcStages, ph = (i / kTcStages) & 1;
      const uint64_t nidx = idx + (uint64_t)kTcProducers * 32;
      const float4 nxt = (i + kTcProducers < nbatch && nidx < rec_end) ? __ldg(reinterpret_cast<const float4*>(recs + nidx))
                                                                          : make_float4(0.f, 0.f, 0.f, 0.f);
      // unpack: origin inside the tile, first plane of the window, 0.18 fixed-point fractions
      const uint32_t glo = __float_as_uint(cur.x), ghi = __float_as_uint(cur.y);
      const uint64_t geo = ((uint64_t)ghi << 32) | glo;
      const int i0 = glo & 7, j0 = (glo >> 3) & 7, p0 = (glo >> 6) & 15;
      const float fu = (float)((glo >> 10) & 0x3ffffu) * (1.0f / 262144.0f);
      const float fv = (float)((uint32_t)(geo >> 28) & 0x3ffffu) * (1.0f / 262144.0f);
      const float fw = (float)(ghi >> 14) * (1.0f / 262144.0f);
      const float xu0 = fmaf(fu - (float)i0, two_over_w, -1.0f), xv0 = fmaf(fv - (float)j0, two_over_w, -1.0f);
      const float xw0 = fmaf(fw - (float)p0, two_over_w, -1.0f);
      float ku[8], kv[8], kw[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float c0 = (float)r * two_over_w, c1 = (float)(r + 8) * two_over_w;
        ku[r] = es_tap(xu0 + (r < i0 ? c1 : c0), bw2);
        kv[r] = es_tap(xv0 + (r < j0 ? c1 : c0), bw2);
        kw[r] = wstack ? es_tap(xw0 + c0, bw2) : (r == 0 ? 1.0f : 0.0f);
      }
      timed_wait(smem_u32(&sh->s_free[s]), ph ^ 1, t_wait0, relax_ns);
      const uint32_t st = sbase + s * kStageBytes;
      if (!(PROF && (dbg & 16))) {
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        sts32(st + r * 128 + lane * 4, ku[r]);
        sts32(st + 1024 + r * 128 + lane * 4, kv[r]);
      }
      const uint32_t bb = st + kStageF + kgrp * kBGroup + kh * kBLbo + k4 * 4;
#pragma unroll
      for (int pl = 0; pl < 8; ++pl) {
        const float re = cur.z * kw[pl], im = cur.w * kw[pl];
        // column n = 2 pl + c holds the hi part (the raw value: the tensor core drops the low 13 bits), n + 16 the lo part
        const int n0 = 2 * pl, n1 = 2 * pl + 1;
        sts32(bb + (n0 / 8) * kBSbo + (n0 % 8) * 16, re);
        sts32(bb + (n1 / 8) * kBSbo + (n1 % 8) * 16, im);
        sts32(bb + ((n0 + 16) / 8) * kBSbo + (n0 % 8) * 16, tf32_lo(re));
        sts32(bb + ((n1 + 16) / 8) * kBSbo + (n1 % 8) * 16, tf32_lo(im));
      }
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&sh->f_full[s]));
      cur = nxt;
      idx = nidx;
    }
    if (PROF && warp == 9 && lane == 0) {
      prof[blockIdx.x * 16 + 0] = clock64() - t_begin;
      prof[blockIdx.x * 16 + 1] = t_wait0;
    }
  } else if (warp < 4) {
    // ================= builders: thread = row of A =================
    const int cell = 16 * warp + (lane >> 1);
    const uint32_t lomask = (lane & 1) ? 0xffffe000u : 0u;  // odd rows carry x - trunc(x), even rows x itself
    const uint32_t fu_off = (cell >> 3) * 128, fv_off = 1024 + (cell & 7) * 128;
    const uint32_t trow = tbase + ((uint32_t)(warp * 32) << 16);
    for (uint32_t i = 0; i < nbatch; ++i) {
      const uint32_t s = i % kTcStages, ph = (i / kTcStages) & 1;
      timed_wait(smem_u32(&sh->f_full[s]), ph, t_wait0);
      const long long tb0 = PROF ? clock64() : 0;
      uint32_t a[32];
      const float4* fu4 = reinterpret_cast<const float4*>(smem + s * kStageBytes + fu_off);
      const float4* fv4 = reinterpret_cast<const float4*>(smem + s * kStageBytes + fv_off);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float4 x, y;
        if (PROF && (dbg & 8)) {
          x = make_float4(1.f + q, 2.f, 3.f, 4.f + lane);
          y = make_float4(0.5f, 0.25f + q, 0.125f, 1.f);
        } else {
          x = fu4[q];
          y = fv4[q];
        }
        const float p0 = x.x * y.x, p1 = x.y * y.y, p2 = x.z * y.z, p3 = x.w * y.w;
        a[4 * q + 0] = __float_as_uint(p0 - __uint_as_float(__float_as_uint(p0) & lomask));
        a[4 * q + 1] = __float_as_uint(p1 - __uint_as_float(__float_as_uint(p1) & lomask));
        a[4 * q + 2] = __float_as_uint(p2 - __uint_as_float(__float_as_uint(p2) & lomask));
        a[4 * q + 3] = __float_as_uint(p3 - __uint_as_float(__float_as_uint(p3) & lomask));
      }
      if (PROF) t_alu += clock64() - tb0;
      if (!(PROF && (dbg & 1))) tmem_st32(trow + s * 32, a);
      const long long tw = PROF ? clock64() : 0;
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      if (PROF) t_wait1 += clock64() - tw;
      fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&sh->a_full[s]));
    }
    if (PROF && warp == 0 && lane == 0) {
      prof[blockIdx.x * 16 + 2] = clock64() - t_begin;
      prof[blockIdx.x * 16 + 3] = t_wait0;
      prof[blockIdx.x * 16 + 4] = t_wait1;
      prof[blockIdx.x * 16 + 13] = t_alu;
    }
  } else if (warp == 8) {
    // ================= issuer: the whole warp runs the loop (uniform values), one elected lane issues =================
    uint32_t j = j0, jn = 0;               // current job, jobs finished
    uint32_t job_end = gbeg + (jobs[j0].ngroups & ~kJobStartsItem);
    uint32_t next_ngroups = j0 + 1 < j1 ? (jobs[j0 + 1].ngroups & ~kJobStartsItem) : 0;
    uint32_t accumulate = 0;                // 0: the next mma starts a window (overwrites the accumulator)
    uint32_t acc = 0;
    timed_wait(smem_u32(&sh->acc_free[0]), 1, t_wait1);
    uint32_t gi = gbeg;
    const uint64_t bdesc0 = make_bdesc(sbase + kStageF);
    for (uint32_t i = 0; i < nbatch; ++i) {
      const uint32_t s = i % kTcStages, ph = (i / kTcStages) & 1;
      timed_wait(smem_u32(&sh->f_full[s]), ph, t_wait0);
      timed_wait(smem_u32(&sh->a_full[s]), ph, t_wait0);
      fence_after();
      const long long tm0 = PROF ? clock64() : 0;
      const uint32_t a_base = tbase + s * 32;
      const uint64_t b_base = bdesc0 + (uint64_t)((s * kStageBytes) >> 4);
      if (gend - gi >= 4 && job_end - gi >= 4) {
        // the whole batch belongs to the current window (3 of 4 batches): four back-to-back mma, operand
        // addresses a constant apart
        if (elect_one()) {
          const uint32_t dcol = tbase + 256 + acc * 32;
          if (!(PROF && (dbg & 4))) {
            tc_mma_ts(dcol, a_base, b_base, kIdesc, accumulate);
            tc_mma_ts(dcol, a_base + 8, b_base + (kBGroup >> 4), kIdesc, 1u);
            tc_mma_ts(dcol, a_base + 16, b_base + 2 * (kBGroup >> 4), kIdesc, 1u);
            tc_mma_ts(dcol, a_base + 24, b_base + 3 * (kBGroup >> 4), kIdesc, 1u);
          }
          tc_commit(smem_u32(&sh->s_free[s]));
        }
        accumulate = 1;
        gi += 4;
        if (PROF) t_mma += clock64() - tm0;
        continue;
      }
      for (uint32_t q = 0; q < 4; ++q, ++gi) {
        if (gi >= gend) break;
        if (gi == job_end) {  // the window is complete: hand it to the epilogue, take the next accumulator
          if (elect_one()) tc_commit(smem_u32(&sh->acc_full[acc]));
          ++j;
          ++jn;
          job_end += next_ngroups;
          next_ngroups = j + 1 < j1 ? (jobs[j + 1].ngroups & ~kJobStartsItem) : 0;
          acc = jn % kTcAccs;
          timed_wait(smem_u32(&sh->acc_free[acc]), ((jn / kTcAccs) & 1) ^ 1, t_wait1);
          fence_after();
          accumulate = 0;
        }
        if (!(PROF && (dbg & 4)) && elect_one())
          tc_mma_ts(tbase + 256 + acc * 32, a_base + q * 8, b_base + q * (kBGroup >> 4), kIdesc, accumulate);
        accumulate = 1;
      }
      if (PROF) t_mma += clock64() - tm0;
      if (elect_one()) tc_commit(smem_u32(&sh->s_free[s]));
    }
    if (elect_one()) tc_commit(smem_u32(&sh->acc_full[acc]));
    __syncwarp();
    if (PROF && lane == 0) {
      prof[blockIdx.x * 16 + 5] = clock64() - t_begin;
      prof[blockIdx.x * 16 + 6] = t_wait0;
      prof[blockIdx.x * 16 + 7] = t_wait1;
      prof[blockIdx.x * 16 + 11] = nbatch;
      prof[blockIdx.x * 16 + 12] = j1 - j0;
      prof[blockIdx.x * 16 + 14] = t_mma;
    }
  } else {
    // ================= epilogue: thread = row of the accumulator =================
    const int q = warp - 4;
    const int cell = 16 * q + (lane >> 1), hl = lane & 1;
    const int cu = cell >> 3, cv = cell & 7;
    const uint32_t trow = tbase + ((uint32_t)(q * 32) << 16) + 256;
    const int snap = 9 - g.W;
    const int64_t plane_cells = (int64_t)g.nu * g.nv;
    uint4 jnext = __ldg(reinterpret_cast<const uint4*>(jobs + j0));
    for (uint32_t j = j0; j < j1; ++j) {
      const uint32_t jn = j - j0, acc = jn % kTcAccs;
      const uint4 jw = jnext;  // {group_begin, ngroups, tu | tv << 16, win | lu << 16 | lv << 24}
      if (j + 1 < j1) jnext = __ldg(reinterpret_cast<const uint4*>(jobs + j + 1));
      timed_wait(smem_u32(&sh->acc_full[acc]), (jn / kTcAccs) & 1, t_wait0, relax_ns);
      fence_after();
      uint32_t d[32];
      if (!(PROF && (dbg & 2))) tmem_ld32(trow + acc * 32, d);
      else for (int z = 0; z < 32; ++z) d[z] = 0;
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&sh->acc_free[acc]));
      // columns n and n + 16 are the hi and lo parts of B; rows 2c and 2c + 1 those of A.  The even row takes
      // planes 0 .. 3 of the cell, the odd row planes 4 .. 7.
      float out[8];
#pragma unroll
      for (int m = 0; m < 8; ++m) {
        const float lo4 = __uint_as_float(d[m]) + __uint_as_float(d[m + 16]);          // planes 0 .. 3
        const float hi4 = __uint_as_float(d[m + 8]) + __uint_as_float(d[m + 24]);      // planes 4 .. 7
        const float mine = hl ? hi4 : lo4, send = hl ? lo4 : hi4;
        out[m] = mine + __shfl_xor_sync(0xffffffffu, send, 1);
      }
      const int wi = (int)((jw.w >> 16) & 0xffu) * snap, wj = (int)(jw.w >> 24) * snap;
      const int ci = wi + ((cu - wi) & 7), cj = wj + ((cv - wj) & 7);
      int gi = (int)(jw.z & 0xffffu) * kTile - kHalo + g.nu / 2 + ci;
      int gj = (int)(jw.z >> 16) * kTile - kHalo + g.nv / 2 + cj;
      gi -= gi >= g.nu ? g.nu : 0;
      gi -= gi >= g.nu ? g.nu : 0;
      gj -= gj >= g.nv ? g.nv : 0;
      gj -= gj >= g.nv ? g.nv : 0;
      const int P0 = (int)(jw.w & 0xffffu) * g.win_step + 4 * hl;
      float* cellp = grid + 2 * ((int64_t)gi * g.nv + gj);
#pragma unroll
      for (int pl = 0; pl < 4; ++pl) {
        const int plane = P0 + pl, slot = plane - plane_base;
        const float re = out[2 * pl], im = out[2 * pl + 1];
        if (plane < g.nplanes && slot >= 0 && slot < nslots && (re != 0.f || im != 0.f))
          asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(cellp + 2 * (int64_t)slot * plane_cells), "f"(re), "f"(im) : "memory");
      }
    }
    if (PROF && warp == 4 && lane == 0) {
      prof[blockIdx.x * 16 + 8] = clock64() - t_begin;
      prof[blockIdx.x * 16 + 9] = t_wait0;
    }
  }
  // ---- teardown ----
  fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(512u) : "memory");
}

}  // namespace

cip_status run_gridder_tc(cip_ctx* ctx, int slot, cudaStream_t stream, const GridGeom& g, int win_first, int win_last, float* grid,
                          int plane_base, int nslots) {
  const BinWork& w = ctx->work[slot];
  static const bool prof = getenv("CIP_TC_PROF") != nullptr;
  static const uint32_t relax_ns = getenv("CIP_TC_RELAX_NS") ? (uint32_t)atoi(getenv("CIP_TC_RELAX_NS")) : 100u;
  static const int dbg = getenv("CIP_TC_DBG") ? atoi(getenv("CIP_TC_DBG")) : 0;  // development: switch parts off (results wrong)
  static const int job_cost = getenv("CIP_TC_JOB_COST") ? atoi(getenv("CIP_TC_JOB_COST")) : kTcJobCost;
  if (prof) {
    long long* d_prof = nullptr;
    const size_t bytes = sizeof(long long) * 16 * ctx->sm_count;
    CIP_CHECK_CUDA(cudaMalloc(&d_prof, bytes));
    CIP_CHECK_CUDA(cudaMemsetAsync(d_prof, 0, bytes, stream));
    CIP_CHECK_CUDA(cudaFuncSetAttribute(grid_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmem));
    grid_tc_kernel<true><<<ctx->sm_count, kTcThreads, kTcSmem, stream>>>(w.jobs.as<Job>(), w.win_job_start.as<uint32_t>(),
                                                                         w.hdr.as<BinHeader>(), w.recs.as<RecP>(), g, grid, win_first,
                                                                         win_last, plane_base, nslots, job_cost, relax_ns, dbg, d_prof);
    std::vector<long long> h(16 * ctx->sm_count);
    CIP_CHECK_CUDA(cudaStreamSynchronize(stream));
    CIP_CHECK_CUDA(cudaMemcpy(h.data(), d_prof, bytes, cudaMemcpyDeviceToHost));
    cudaFree(d_prof);
    double a[16] = {};
    for (int c = 0; c < ctx->sm_count; ++c)
      for (int k = 0; k < 16; ++k) a[k] += (double)h[16 * c + k] / ctx->sm_count;
    fprintf(stderr, "[tc prof] per CTA: batches %.0f jobs %.0f | producer total %.0f wait(s_free) %.0f | builder total %.0f wait(f_full) %.0f "
                    "wait::st %.0f alu %.0f | issuer total %.0f wait(full) %.0f wait(acc_free) %.0f mma-loop %.0f | epilogue total %.0f wait(acc_full) %.0f cycles\n",
            a[11], a[12], a[0], a[1], a[2], a[3], a[4], a[13], a[5], a[6], a[7], a[14], a[8], a[9]);
  } else {
    CIP_CHECK_CUDA(cudaFuncSetAttribute(grid_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmem));
    grid_tc_kernel<false><<<ctx->sm_count, kTcThreads, kTcSmem, stream>>>(w.jobs.as<Job>(), w.win_job_start.as<uint32_t>(),
                                                                          w.hdr.as<BinHeader>(), w.recs.as<RecP>(), g, grid, win_first,
                                                                          win_last, plane_base, nslots, job_cost, relax_ns, dbg, nullptr);
  }
  ctx->launches++;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

}  // namespace cip
