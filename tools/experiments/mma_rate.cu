// Issue rate of the legacy tensor path (mma.sync) on one SM sub-partition, for the shapes a tensor-core
// gridder would use: m16n8k8 / m16n8k4 tf32 and m16n8k16 bf16, 8 independent accumulator tiles per warp (the 64 x 16
// register window).  Prints cycles per mma per sub-partition for 1..4 warps per sub-partition.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int KIND>
__global__ void rate_kernel(float* out, int iters, long long* cycles) {
  float c[8][4];
  for (int i = 0; i < 8; ++i)
    for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
  uint32_t a[4] = {threadIdx.x * 3u + 1u, threadIdx.x * 5u, threadIdx.x + 7u, threadIdx.x * 11u};
  uint32_t b[2] = {threadIdx.x * 13u, threadIdx.x * 17u + 3u};
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int rep = 0; rep < 3; ++rep)
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        if (KIND == 0)
          asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                       : "+f"(c[t][0]), "+f"(c[t][1]), "+f"(c[t][2]), "+f"(c[t][3])
                       : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
        else if (KIND == 2)
          asm volatile("mma.sync.aligned.m16n8k4.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                       : "+f"(c[t][0]), "+f"(c[t][1]), "+f"(c[t][2]), "+f"(c[t][3])
                       : "r"(a[0]), "r"(a[1]), "r"(b[0]));
        else
          asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                       : "+f"(c[t][0]), "+f"(c[t][1]), "+f"(c[t][2]), "+f"(c[t][3])
                       : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
      }
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < 8; ++i)
    for (int j = 0; j < 4; ++j) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

int main() {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4);
  cudaMallocManaged(&cyc, 8);
  const int iters = 2000;
  for (int kind = 0; kind < 3; ++kind)
    for (int wps = 1; wps <= 4; ++wps) {  // warps per sub-partition
      for (int r = 0; r < 2; ++r) {
        if (kind == 0) rate_kernel<0><<<148, 128 * wps>>>(out, iters, cyc);
        else if (kind == 2) rate_kernel<2><<<148, 128 * wps>>>(out, iters, cyc);
        else rate_kernel<1><<<148, 128 * wps>>>(out, iters, cyc);
        cudaDeviceSynchronize();
      }
      double per = (double)*cyc / ((double)iters * 24 * wps);
      printf("%s warps/SMSP %d: %.2f cycles per mma per SMSP (%s)\n", kind == 2 ? "m16n8k4 tf32" : kind ? "m16n8k16 bf16" : "m16n8k8 tf32", wps, per,
             cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
