// cuFFT 1-D C2C along contiguous rows of a square plane, whole plane in one exec against blocks of rows:
// how much the block-wise form (which lets empty row blocks be skipped) costs.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fft_blocks fft_blocks.cu -lcufft && ./fft_blocks
#include <cstdio>
#include <cuda_runtime.h>
#include <cufft.h>

int main() {
  for (int n : {6144, 12288, 24576}) {
    cufftComplex* d;
    cudaMalloc(&d, sizeof(cufftComplex) * (size_t)n * n);
    cudaMemset(d, 0, sizeof(cufftComplex) * (size_t)n * n);
    for (int nb : {1, 8, 16, 32, 64}) {
      const int batch = n / nb;
      cufftHandle h;
      int nn[1] = {n};
      size_t work = 0;
      cufftCreate(&h);
      if (cufftMakePlanMany(h, 1, nn, nn, 1, n, nn, 1, n, CUFFT_C2C, batch, &work) != CUFFT_SUCCESS) { printf("plan failed\n"); continue; }
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        for (int it = 0; it < 4; ++it)
          for (int b = 0; b < nb; ++b) cufftExecC2C(h, d + (size_t)b * batch * n, d + (size_t)b * batch * n, CUFFT_INVERSE);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
      }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      printf("n %5d  blocks %2d (batch %5d)  %.3f ms per plane  work %.1f MB\n", n, nb, batch, ms / 4, work / 1e6);
      cufftDestroy(h);
    }
    cudaFree(d);
  }
  return 0;
}
