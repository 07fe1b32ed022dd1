// Micro-benchmark: shared-memory wavefronts of 128-bit / 64-bit loads for the lane -> address patterns
// considered for the gridder's operand stage.  Run under
//   ncu --metrics l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,smsp__inst_executed_op_shared_ld.sum
// and divide: wavefronts per LDS instruction of each pattern (one kernel launch per pattern).
#include <cstdio>
#include <cuda_runtime.h>

template <int P>
__device__ __forceinline__ unsigned lane_offset(int lane) {
  const int a = lane & 7;
  switch (P) {
    case 0: return 0;                                                        // all lanes one address
    case 1: return ((lane >> 3) & 1) * 32;                                   // 2 addresses per half-warp (ph = bit 3)
    case 2: return (a * 2 + (lane >> 4)) * 16;                               // stride-32B chunks, quarters 0/1 identical
    case 3: return ((lane >> 4) * 8 + a) * 16;                               // contiguous 128 B per quarter, quarters 0/1 identical
    case 4: return (((lane >> 3) & 1) * 8 + a) * 16;                         // contiguous, quarters 0/2 identical
    case 5: return lane * 16;                                                // 32 distinct chunks
    case 6: return (lane >> 4) * 32;                                         // one address per half-warp
    case 7: return (lane >> 3) * 32;                                         // one address per quarter-warp
    case 8: return ((lane >> 1) & 7) * 16;                                   // lane pairs share, 8 chunks per half-warp contiguous
    case 9: return (lane >> 1) * 16;                                         // lane pairs share, 16 chunks per warp
    case 10: return (lane & 15) * 16;                                        // halves identical, 16 chunks each (256 B)
    case 11: return (lane & 3) * 16 + (lane >> 4) * 64;                      // 4 chunks per quarter, quarters 0/1 identical
  }
  return 0;
}

template <int P, int BYTES>
__global__ void k(float* out, int iters) {
  __shared__ __align__(16) unsigned char sm[8192];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = (float)i;
  __syncthreads();
  unsigned base = (unsigned)__cvta_generic_to_shared(sm) + lane_offset<P>(threadIdx.x & 31) * (BYTES == 16 ? 1 : 1);
  float acc = 0.f;
  for (int i = 0; i < iters; ++i) {
    unsigned addr = base + (i & 7) * 528;
    if (BYTES == 16) {
      float4 v;
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
      acc += v.x + v.y + v.z + v.w;
    } else {
      float2 v;
      asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
      acc += v.x + v.y;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int P>
void run(float* out) {
  k<P, 16><<<148, 32>>>(out, 4096);
  k<P, 8><<<148, 32>>>(out, 4096);
}

int main() {
  float* out;
  cudaMalloc(&out, 148 * 32 * 4);
  run<0>(out); run<1>(out); run<2>(out); run<3>(out); run<4>(out); run<5>(out);
  run<6>(out); run<7>(out); run<8>(out); run<9>(out); run<10>(out); run<11>(out);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
