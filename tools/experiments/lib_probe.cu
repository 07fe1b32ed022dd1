// Development probe: cip_invert with cudaHostAlloc'ed inputs from a plain C++ process (no Python, no torch).
//   nvcc -gencode arch=compute_100a,code=sm_100a -I../../include -o lib_probe lib_probe.cu -L<libdir> -lcip_b200
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "cip.h"
extern "C" int cip_debug_overlap(cip_ctx*, const void*, size_t, const void*, size_t, int);
int main(int argc, char** argv) {
  const int64_t nrow = argc > 1 ? atoll(argv[1]) : 1317888, nchan = 64;
  double *uvw, *freq; float *vis, *wgt;
  cudaHostAlloc(&uvw, nrow * 24, 0); cudaHostAlloc(&freq, nchan * 8, 0);
  cudaHostAlloc(&vis, nrow * nchan * 8, 0); cudaHostAlloc(&wgt, nrow * nchan * 4, 0);
  srand(1);
  for (int64_t r = 0; r < nrow; ++r) {
    double a = 6.2831853 * rand() / RAND_MAX, rad = 30.0 * exp(7.8 * rand() / RAND_MAX);
    uvw[3 * r] = rad * cos(a); uvw[3 * r + 1] = rad * sin(a); uvw[3 * r + 2] = 0.01 * rad * (rand() / (double)RAND_MAX - 0.5);
  }
  for (int c = 0; c < nchan; ++c) freq[c] = 1.3e9 + 1e6 * c;
  for (int64_t k = 0; k < nrow * nchan; ++k) { vis[2 * k] = 1.f; vis[2 * k + 1] = 0.5f; wgt[k] = 1.f; }
  cip_ctx* ctx; if (cip_ctx_create(0, &ctx)) { printf("%s\n", cip_last_error()); return 1; }
  float* img; cudaHostAlloc(&img, 4096 * 4096 * 4, 0);
  cip_invert_args a{}; a.uvw = uvw; a.nrow = nrow; a.freq = freq; a.nchan = nchan; a.vis = vis; a.wgt = wgt;
  a.npix_x = a.npix_y = 4096; a.pixsize_x = a.pixsize_y = 0.4 / (75000.0 * 1.364e9 / 299792458.0); a.epsilon = 1e-4;
  a.do_wstacking = 1; a.memspace = 0;
  if (getenv("PROBE_OVERLAP")) {
    int when = atoi(getenv("PROBE_OVERLAP"));
    if (when == 2) { cip_report rep{}; cip_invert(ctx, &a, img, nullptr, nullptr, &rep); }
    for (int v = 0; v < 2; ++v) for (int k = 0; k < 2; ++k) cip_debug_overlap(ctx, wgt, nrow * nchan * 4, vis, nrow * nchan * 8, v);
    return 0;
  }
  for (int it = 0; it < 3; ++it) {
    cip_report rep{};
    fprintf(stderr, "--- iteration %d\n", it);
    if (cip_invert(ctx, &a, img, nullptr, nullptr, &rep)) { printf("%s\n", cip_last_error()); return 1; }
    printf("it %d total %.2f ms h2d %.2f bin %.2f grid %.2f\n", it, rep.ms_total, rep.ms_h2d, rep.ms_bin, rep.ms_grid);
  }
  return 0;
}
