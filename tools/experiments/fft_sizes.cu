// cuFFT 2-D C2C in-place timing for candidate oversampled grid sizes (one plane each).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fft_sizes fft_sizes.cu -lcufft
#include <cstdio>
#include <cuda_runtime.h>
#include <cufft.h>
int main(int argc, char** argv) {
  int sizes[] = {5120, 6144, 7168, 8192, 10240, 12288, 14336, 16384, 20480, 24576, 28672, 32768};
  for (int n : sizes) {
    cufftComplex* d;
    size_t bytes = (size_t)n * n * sizeof(cufftComplex);
    if (cudaMalloc(&d, bytes) != cudaSuccess) { printf("%d alloc failed\n", n); continue; }
    cudaMemset(d, 0, bytes);
    cufftHandle h; size_t work = 0;
    cufftCreate(&h);
    int nn[2] = {n, n};
    if (cufftMakePlanMany(h, 2, nn, nullptr, 1, n * n, nullptr, 1, n * n, CUFFT_C2C, 1, &work) != CUFFT_SUCCESS) { printf("%d plan failed\n", n); cudaFree(d); continue; }
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cufftExecC2C(h, d, d, CUFFT_INVERSE);
    cudaEventRecord(e0);
    for (int i = 0; i < 3; ++i) cufftExecC2C(h, d, d, CUFFT_INVERSE);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
    printf("n=%6d  %8.3f ms  %6.1f ps/cell  work %zu MB  %.0f GB/s (4 x plane bytes)\n", n, ms, ms * 1e9 / ((double)n * n), work >> 20, 4.0 * bytes / ms / 1e6);
    cufftDestroy(h); cudaFree(d);
  }
  return 0;
}
