// tcgen05 / TMEM building blocks of a tensor-core gridder, checked and timed in isolation (sm_100a):
//
//   check: D[M x N] = sum over G k-groups of A[M x 8] B[8 x N], kind::tf32, accumulators in TMEM, read back with
//          tcgen05.ld and compared with a host product of the TF32-truncated operands.  Two operand routes:
//            TS  A written to TMEM by the threads (tcgen05.st.32x32b.x8, thread = row), B in shared memory
//            SS  A and B in shared memory
//          Shared-memory operands are MN-major (the MN index is contiguous), no swizzle: a visibility (k) owns one
//          16-byte piece of every core matrix, which is what a "one visibility per lane" producer writes.
//   trunc: whether the tensor core truncates or rounds the 13 low mantissa bits of a float32 operand.
//   rate:  cycles per tcgen05.mma for back-to-back issues of one thread (M = 64 / 128, N = 16 / 32 / 64, TS and SS).
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tcgen05_probe tcgen05_probe.cu && ./tcgen05_probe
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess) {                                                               \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);      \
      exit(2);                                                                             \
    }                                                                                      \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[tmem] B[smem]
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
// D[tmem] (+)= A[smem] B[smem]
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
               "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Instruction descriptor of kind::tf32 with float32 accumulation (bit layout: cute/arch/mma_sm100_desc.hpp).
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// Shared-memory matrix descriptor, no swizzle, MN-major: element (mn, k) of a k-group of 8 sits at
//   start + (mn / 4) * sbo + (k % 8) * 16 + (mn % 4) * 4   bytes;   the next k-group at + lbo.
__device__ __forceinline__ uint64_t make_sdesc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}

constexpr int kMaxG = 4;

// A: (G*8, M) row k, column m (k-major list of columns of A);  B: (G*8, N);  D: (M, N) row-major.
template <int M, int N, bool TS>
__global__ void __launch_bounds__(128) check_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D,
                                                    int G, uint32_t idesc, int reps, long long* cycles, int variant) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) unsigned long long s_bar;
  const int tid = threadIdx.x, warp = tid >> 5;
  float* sB = reinterpret_cast<float*>(smem);                        // G groups of N * 8 floats
  float* sA = reinterpret_cast<float*>(smem + kMaxG * 256 * 8 * 4);  // G groups of M * 8 floats (SS only)
  if (warp == 0) tmem_alloc(smem_u32(&s_tmem), 512);
  if (tid == 0) mbar_init(smem_u32(&s_bar), 1);
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tbase = s_tmem;
  const uint32_t dcol = 0, acol = 256;
  // operands -> shared memory in the canonical MN-major layout
  for (int i = tid; i < G * 8 * N; i += 128) {
    const int k = i / N, n = i % N, g = k / 8;
    if (variant & 8) sB[g * N * 8 + (n / 8) * 64 + ((k % 8) / 4) * 32 + (n % 8) * 4 + (k % 4)] = B[(size_t)k * N + n];  // K-major: 8 n-rows x 16 B core matrices, the two k-chunks 128 B apart, n-groups 256 B apart
    else sB[g * N * 8 + (n / 4) * 32 + (k % 8) * 4 + (n % 4)] = B[(size_t)k * N + n];
  }
  if (variant & 1) {  // sentinel in the accumulator columns
    uint32_t one[8];
    for (int k = 0; k < 8; ++k) one[k] = __float_as_uint(1.0f + k);
    for (int c = 0; c < N; c += 8) tmem_st8(tbase + ((uint32_t)(warp * 32) << 16) + dcol + c, one);
    tmem_wait_st();
  }
  if (!TS)
    for (int i = tid; i < G * 8 * M; i += 128) {
      const int k = i / M, m = i % M, g = k / 8;
      if (variant & 16) sA[g * M * 8 + (m / 4) * 32 + (k % 8) * 4 + (m % 4)] = A[(size_t)k * M + m];  // MN-major
      else sA[g * M * 8 + (m / 8) * 64 + ((k % 8) / 4) * 32 + (m % 8) * 4 + (k % 4)] = A[(size_t)k * M + m];  // K-major
    }
  if (TS && tid < M) {  // thread = row of A: 8 columns per k-group
    for (int g = 0; g < G; ++g) {
      uint32_t v[8];
      for (int k = 0; k < 8; ++k) v[k] = __float_as_uint(A[(size_t)(g * 8 + k) * M + tid]);
      tmem_st8(tbase + ((uint32_t)(warp * 32) << 16) + acol + 8 * g, v);
    }
    tmem_wait_st();
  }
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  long long t0 = 0, t1 = 0;
  if (tid == 0) {
    t0 = clock64();
    for (int r = 0; r < reps; ++r)
      for (int g = 0; g < G; ++g) {
        uint64_t bd = make_sdesc(smem_u32(sB + g * N * 8), 128u * (N / 4), 128u);
        if (variant & 4) bd = make_sdesc(smem_u32(sB + g * N * 8), 128u, 128u * (N / 4));
        if (variant & 8) bd = make_sdesc(smem_u32(sB + g * N * 8), 128u, 256u);
        if (variant & 2) continue;
        if (TS) {
          mma_ts(tbase + dcol + ((variant & 32) ? (g & 1) * N : 0), tbase + acol + 8 * g, bd, idesc, ((variant & 32) ? (r | (g >> 1)) : (r | g)) != 0);
        } else {
          const uint64_t ad = (variant & 16) ? make_sdesc(smem_u32(sA + g * M * 8), 128u * (M / 4), 128u)
                                             : make_sdesc(smem_u32(sA + g * M * 8), 128u, 256u);
          mma_ss(tbase + dcol, ad, bd, idesc, (r | g) != 0);
        }
      }
    commit(smem_u32(&s_bar));
    mbar_wait(smem_u32(&s_bar), 0);
    t1 = clock64();
    if (cycles) *cycles = t1 - t0;
  }
  __syncthreads();
  fence_after();
  // read D back: the M = 128 layout puts row m in lane m; the M = 64 layout puts row m in lane (m % 16) + 32 (m / 16)
  for (int c0 = 0; c0 < N; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(tbase + ((uint32_t)(warp * 32) << 16) + dcol + c0, v);
    tmem_wait_ld();
    const int lane = tid & 31;
    int row = -1;
    if (M == 128) row = tid;
    else if (lane < 16) row = 16 * warp + lane;
    if (row >= 0)
      for (int j = 0; j < 16; ++j) D[(size_t)row * N + c0 + j] = __uint_as_float(v[j]);
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}


__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred)::"memory");
  return pred;
}

// Issue rate with the descriptors prepared outside the loop: warp 0 enters uniformly, one elected lane issues
// 8 mma per iteration (4 k-groups x 2 accumulators when ALT, else one accumulator).
template <int M, int N, bool TS, bool ALT>
__global__ void __launch_bounds__(128) rate_kernel(int reps, long long* cycles, float* sink, int variant) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ __align__(8) unsigned long long s_bar2;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) mbar_init(smem_u32(&s_bar2), 1);
  for (int i = tid; i < (4 * 256 * 8 + 4 * 128 * 8); i += 128) reinterpret_cast<float*>(smem)[i] = 0.001f * (i & 63);
  if (warp == 0) tmem_alloc(smem_u32(&s_tmem), 512);
  if (tid == 0) mbar_init(smem_u32(&s_bar), 1);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tbase = s_tmem;
  const uint32_t idesc = make_idesc(M, N, 0, 0);
  if (warp == 0) {
    uint64_t bd[4], ad[4];
    for (int g = 0; g < 4; ++g) {
      bd[g] = (variant & 1) ? make_sdesc(smem_u32(smem + g * 1184), 144u, 288u) : make_sdesc(smem_u32(smem + g * N * 32), 128u, 256u);
      ad[g] = make_sdesc(smem_u32(smem + 4 * 256 * 32 + g * M * 32), 128u, 256u);
    }
    const uint32_t d0 = tbase, d1 = tbase + (ALT ? N : 0), a0 = tbase + 256;
    __syncwarp();
    const long long t0 = clock64();
    if (elect_one()) {
      for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          if (TS) {
            mma_ts(d0, a0 + 8 * g, bd[g], idesc, 1);
            mma_ts(d1, a0 + 32 + 8 * g, bd[g], idesc, 1);
            if ((variant & 2) && g == 3) commit(smem_u32(&s_bar2));
          } else {
            mma_ss(d0, ad[g], bd[g], idesc, 1);
            mma_ss(d1, ad[g], bd[g], idesc, 1);
          }
        }
      }
      commit(smem_u32(&s_bar));
    }
    __syncwarp();
    const long long t1 = clock64();
    mbar_wait(smem_u32(&s_bar), 0);
    const long long t2 = clock64();
    if (tid == 0) {
      cycles[0] = t1 - t0;
      cycles[1] = t2 - t0;
    }
  }
  __syncthreads();
  fence_after();
  uint32_t v[16];
  tmem_ld16(tbase + ((uint32_t)(warp * 32) << 16), v);
  tmem_wait_ld();
  sink[tid] = __uint_as_float(v[0]);
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

template <int M, int N, bool TS, bool ALT>
static void run_rate(int reps, int variant = 0) {
  long long* cyc;
  float* sink;
  CK(cudaMallocManaged(&cyc, 16));
  CK(cudaMalloc(&sink, 128 * 4));
  const int smem = 4 * 256 * 32 + 4 * 128 * 32;
  auto kern = rate_kernel<M, N, TS, ALT>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (int it = 0; it < 2; ++it) {
    kern<<<1, 128, smem>>>(reps, cyc, sink, variant);
    CK(cudaDeviceSynchronize());
  }
  printf("variant %d ", variant);
  printf("lean rate %s M=%3d N=%3d %s: %5d mma: issue %.2f cycles per mma, complete %.2f cycles per mma\n", TS ? "TS" : "SS", M, N,
         ALT ? "2 accumulators" : "1 accumulator ", reps * 8, (double)cyc[0] / (reps * 8), (double)cyc[1] / (reps * 8));
  cudaFree(cyc);
  cudaFree(sink);
}

static float tf32_trunc(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  u &= 0xffffe000u;
  memcpy(&x, &u, 4);
  return x;
}
static float tf32_round(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  u = (u + 0x1000u) & 0xffffe000u;
  memcpy(&x, &u, 4);
  return x;
}

template <int M, int N, bool TS>
static void run_check(int G, int reps, bool verbose, int variant = 0) {
  const int K = G * 8;
  std::vector<float> A((size_t)K * M), B((size_t)K * N), D((size_t)M * N, -1.f);
  uint32_t s = 12345u + M * 7 + N * 3 + TS;
  auto rnd = [&]() {
    s = s * 1664525u + 1013904223u;
    return (float)((s >> 8) & 0xffff) / 65536.0f - 0.5f;
  };
  for (auto& x : A) x = rnd();
  for (auto& x : B) x = rnd();
  float *dA, *dB, *dD;
  long long* cyc;
  CK(cudaMalloc(&dA, A.size() * 4));
  CK(cudaMalloc(&dB, B.size() * 4));
  CK(cudaMalloc(&dD, D.size() * 4));
  CK(cudaMallocManaged(&cyc, 8));
  CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0xff, D.size() * 4));
  const uint32_t idesc = make_idesc(M, N, (variant & 16) ? 1 : 0, (variant & 8) ? 0 : 1);
  const int smem = kMaxG * 256 * 8 * 4 + kMaxG * 128 * 8 * 4;
  auto kern = check_kernel<M, N, TS>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<1, 128, smem>>>(dA, dB, dD, G, idesc, reps, cyc, variant);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("%s M=%d N=%d G=%d: kernel failed: %s\n", TS ? "TS" : "SS", M, N, G, cudaGetErrorString(e));
    exit(3);
  }
  if (reps == 1) {
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double err_t = 0, err_r = 0, nrm = 0;
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) {
        double rt = 0, rr = 0;
        for (int k = 0; k < K; ++k) {
          rt += (double)tf32_trunc(A[(size_t)k * M + m]) * tf32_trunc(B[(size_t)k * N + n]);
          rr += (double)tf32_round(A[(size_t)k * M + m]) * tf32_round(B[(size_t)k * N + n]);
        }
        const double d = D[(size_t)m * N + n];
        err_t += (d - rt) * (d - rt);
        err_r += (d - rr) * (d - rr);
        nrm += rt * rt;
      }
    printf("variant %2d ", variant);
    printf("check %s M=%3d N=%3d K=%2d: rel err vs truncated operands %.3e, vs rounded operands %.3e  %s\n", TS ? "TS" : "SS", M,
           N, K, sqrt(err_t / nrm), sqrt(err_r / nrm), sqrt(err_t / nrm) < 1e-5 || sqrt(err_r / nrm) < 1e-5 ? "OK" : "MISMATCH");
    if (verbose && !(sqrt(err_t / nrm) < 1e-5))
      for (int m = 0; m < 4; ++m) {
        for (int n = 0; n < 8; ++n) printf(" %9.5f", D[(size_t)m * N + n]);
        printf("\n");
      }
  } else {
    printf("variant %2d ", variant);
    printf("rate  %s M=%3d N=%3d: %d back-to-back mma in %lld cycles = %.2f cycles per mma (K = 8)\n", TS ? "TS" : "SS", M, N,
           reps * G, *cyc, (double)*cyc / (reps * G));
  }
  cudaFree(dA);
  cudaFree(dB);
  cudaFree(dD);
  cudaFree(cyc);
}

int main(int argc, char** argv) {
  // variant bits: 1 sentinel in D, 2 skip the mma, 4 MN-major B with LBO/SBO swapped, 8 K-major B (the layout that
  // works; MN-major no-swizzle operands gave all-zero products on B200 / CUDA 12.9), 16 MN-major A (SS), 32 two
  // alternating accumulators (rate only)
  const char* mode = argc > 1 ? argv[1] : "all";
  const bool all = !strcmp(mode, "all");
  if (!strcmp(mode, "diag")) {
    for (int v : {1, 3, 0, 4, 8, 9}) run_check<128, 32, true>(1, 1, true, v);
    for (int v : {16, 20, 24, 8}) run_check<64, 16, false>(1, 1, true, v);
  }
  if (all || !strcmp(mode, "check")) {
    run_check<64, 16, false>(1, 1, true, 8);
    run_check<64, 16, false>(4, 1, true, 8);
    run_check<128, 32, false>(4, 1, true, 8);
    run_check<128, 32, true>(1, 1, true, 8);
    run_check<128, 32, true>(4, 1, true, 8);
    run_check<128, 16, true>(4, 1, true, 8);
    run_check<128, 64, true>(4, 1, true, 8);
  }
  if (!strcmp(mode, "pad")) {
    for (int v : {0, 1, 2, 3}) run_rate<128, 32, true, true>(256, v);
  }
  if (all || !strcmp(mode, "lean")) {
    for (int reps : {32, 256}) {
      run_rate<64, 16, false, false>(reps);
      run_rate<64, 16, false, true>(reps);
      run_rate<128, 16, false, true>(reps);
      run_rate<128, 32, false, true>(reps);
      run_rate<128, 16, true, false>(reps);
      run_rate<128, 16, true, true>(reps);
      run_rate<128, 32, true, false>(reps);
      run_rate<128, 32, true, true>(reps);
      run_rate<128, 64, true, true>(reps);
      run_rate<128, 128, true, true>(reps);
      run_rate<128, 256, true, false>(reps);
    }
  }
  if (!strcmp(mode, "rate")) {
    for (int reps : {64, 512})
      for (int v : {8, 40}) {
        run_check<64, 16, false>(4, reps, false, v);
        run_check<128, 16, false>(4, reps, false, v);
        run_check<128, 32, false>(4, reps, false, v);
        run_check<128, 64, false>(4, reps, false, v);
        run_check<128, 16, true>(4, reps, false, v);
        run_check<128, 32, true>(4, reps, false, v);
        run_check<128, 64, true>(4, reps, false, v);
        run_check<128, 128, true>(4, reps, false, v);
      }
  }
  return 0;
}
