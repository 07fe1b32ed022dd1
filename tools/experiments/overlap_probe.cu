// Development probe: do kernels on stream A overlap with big H2D copies on stream B, and what does a
// small D2H / mapped write / memset on A cost while they run?   nvcc -arch=sm_100a -o overlap_probe overlap_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <cuda_runtime.h>
#ifdef WITH_CUFFT
#include <cufft.h>
#endif
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)
static double now() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
__global__ void busy(float* p, int n, int iters) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { float v = p[i]; for (int k = 0; k < iters; ++k) v = v * 1.0001f + 0.5f; p[i] = v; }
}
__global__ void put(const int* src, int* dst) { dst[0] = src[0]; __threadfence_system(); }
int main(int argc, char** argv) {
  int mode = argc > 1 ? atoi(argv[1]) : 0;  // 0: mapped write, 1: cudaMemcpyAsync D2H, 2: + memsetAsync 2 GB, 3: no cross-stream wait
  size_t big = (size_t)1350 << 20, small = (size_t)700 << 20;
  char *hbig, *hsmall, *dbig, *dsmall; float* dwork; int *dflag, *hflag, *hflag_dev; char* dzero;
  CK(cudaHostAlloc(&hbig, big, cudaHostAllocDefault)); CK(cudaHostAlloc(&hsmall, small, cudaHostAllocDefault));
  memset(hbig, 1, big); memset(hsmall, 1, small);
  CK(cudaMalloc(&dbig, big)); CK(cudaMalloc(&dsmall, small)); CK(cudaMalloc(&dwork, 64 << 20)); CK(cudaMalloc(&dflag, 4));
  CK(cudaMalloc(&dzero, (size_t)2 << 30));
  CK(cudaHostAlloc(&hflag, 4096, cudaHostAllocMapped)); CK(cudaHostGetDevicePointer(&hflag_dev, hflag, 0));
  cudaStream_t A, B; CK(cudaStreamCreateWithFlags(&A, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&B, cudaStreamNonBlocking));
  cudaEvent_t ev, cev[8]; CK(cudaEventCreate(&ev)); for (int c = 0; c < 8; ++c) CK(cudaEventCreate(&cev[c]));
#ifdef WITH_CUFFT
  { cufftHandle h; cufftPlan1d(&h, 1024, CUFFT_C2C, 1); }
#endif
  if (mode == 9) {
    int sub = argc > 2 ? atoi(argv[2]) : 0;
    cudaEvent_t sev, e14, e12; CK(cudaEventCreateWithFlags(&sev, (sub & 1) ? cudaEventDefault : cudaEventDisableTiming)); CK(cudaEventCreate(&e14)); CK(cudaEventCreate(&e12));
    for (int rep = 0; rep < 3; ++rep) {
      CK(cudaDeviceSynchronize());
      double t0 = now();
      if (!(sub & 4)) CK(cudaEventRecord(e14, A));
      CK(cudaMemcpyAsync(dsmall, hsmall, small / 2, cudaMemcpyHostToDevice, A));
      CK(cudaEventRecord(sev, A));
      CK(cudaStreamWaitEvent(B, sev, 0));
      for (int c = 0; c < 8; ++c) { CK(cudaMemcpyAsync(dbig + c * (big / 16), hbig + c * (big / 16), big / 16, cudaMemcpyHostToDevice, B)); CK(cudaEventRecord(cev[c], B)); }
      CK(cudaEventSynchronize(sev));
      double t1 = now();
      if (!(sub & 2)) CK(cudaEventRecord(e12, A));
      if (sub & 8) busy<<<(16 << 20) / 256, 256, 0, A>>>(dwork, 16 << 20, 100);
      put<<<1, 32, 0, A>>>(dflag, (int*)dwork);
      CK(cudaStreamSynchronize(A));
      double t2 = now();
      CK(cudaStreamSynchronize(B));
      float a = 0, b = 0; cudaEventElapsedTime(&a, e14, e12); cudaEventElapsedTime(&b, e14, cev[7]);
      printf("mode 9 sub %d rep %d: small done %.2f | kernel done %.2f | device clock: kernel-stream event %.3f, last chunk %.3f\n", sub, rep, t1 - t0, t2 - t0, a, b);
    }
    return 0;
  }
  for (int rep = 0; rep < 3; ++rep) {
    CK(cudaDeviceSynchronize());
    double t0 = now();
    if (mode == 6 || mode == 8) CK(cudaMemsetAsync(dwork, 0, 64 << 20, A));
    if (mode == 7 || mode == 8) { CK(cudaMemcpyAsync(dsmall, hsmall, 31 << 20, cudaMemcpyHostToDevice, A)); CK(cudaMemcpyAsync(dflag, hsmall, 4, cudaMemcpyHostToDevice, A)); }
    CK(cudaMemcpyAsync(dsmall, hsmall, small, cudaMemcpyHostToDevice, A));
    CK(cudaEventRecord(ev, A));
    if (mode != 3) CK(cudaStreamWaitEvent(B, ev, 0));
    for (int c = 0; c < 8; ++c) { CK(cudaMemcpyAsync(dbig + c * (big / 8), hbig + c * (big / 8), big / 8, cudaMemcpyHostToDevice, B)); CK(cudaEventRecord(cev[c], B)); }
    double t1 = now();
    if (mode >= 4) CK(cudaEventSynchronize(ev));
    if (mode == 5) { CK(cudaEventRecord(cev[7], A)); }
    busy<<<(16 << 20) / 256, 256, 0, A>>>(dwork, 16 << 20, 100);
    if (mode == 1) { CK(cudaMemcpyAsync(hflag, dflag, 4, cudaMemcpyDeviceToHost, A)); } else { put<<<1, 1, 0, A>>>(dflag, hflag_dev); }
    CK(cudaStreamSynchronize(A));
    double t2 = now();
    if (mode == 2) CK(cudaMemsetAsync(dzero, 0, (size_t)2 << 30, A));
    for (int c = 0; c < 8; ++c) { CK(cudaStreamWaitEvent(A, cev[c], 0)); busy<<<(16 << 20) / 256, 256, 0, A>>>(dwork, 16 << 20, 2000); }
    CK(cudaStreamSynchronize(A));
    double t3 = now();
    CK(cudaStreamSynchronize(B));
    double t4 = now();
    printf("mode %d rep %d: enqueue %.2f ms | first kernel+fetch done at %.2f | chunk kernels done at %.2f | copies done at %.2f\n", mode, rep, t1 - t0, t2 - t0, t3 - t0, t4 - t0);
  }
  return 0;
}
