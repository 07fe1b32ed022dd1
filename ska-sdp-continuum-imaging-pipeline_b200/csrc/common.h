// Internal declarations shared by the translation units of libcip_b200.so.
#pragma once

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <map>
#include <string>
#include <tuple>
#include <vector>

#include <cuda_runtime.h>
#include <cufft.h>
#include <nvtx3/nvToolsExt.h>

#include "cip.h"

namespace cip {

constexpr double kSpeedOfLight = 299792458.0;
constexpr int kMaxSupport = 16;   // widest kernel in the design table
constexpr int kFastSupport = 8;   // widest kernel of the residue-owner gridder (= residue modulus)
constexpr int kMinSupport = 4;
constexpr int kTile = 8;          // uv tile edge of the gridder, in grid cells
constexpr int kHalo = 8;          // shift that keeps window origins non-negative (>= kMaxSupport/2)
constexpr int kSub = kTile + kFastSupport - 1;  // sub-grid edge held in shared memory (15, odd on purpose)
constexpr int kWin = 8;           // w planes of one gridding window (= planes per consumer thread)
constexpr int kBatch = 32;        // visibilities per producer/consumer hand-over
constexpr int kGroup = 8;          // records per k-group of the tensor-core gridder: every window job is padded to a multiple
constexpr int kItemGroups = 256;  // k-groups per work item of the register-window gridders (2048 records, measured in round 1)
constexpr uint32_t kJobStartsItem = 0x80000000u;  // bit 31 of Job::ngroups: the job starts a work item
constexpr int kJobGroups = 512;   // k-groups (of 8 records) per gridding job at most: bounds the length of one accumulation
constexpr int kMaxChunks = 8;     // row chunks of the host -> device visibility stream

void set_error(const char* fmt, ...);
// CIP_TRACE=1: host wall-clock marks on stderr (development aid; no device synchronisation).
void trace_mark(const char* label);
double now_ms();  // host monotonic clock

#define CIP_CHECK_CUDA(expr)                                                                       \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) {                                                                       \
      ::cip::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return _e == cudaErrorMemoryAllocation ? CIP_ENOMEM : CIP_ECUDA;                             \
    }                                                                                              \
  } while (0)

#define CIP_CHECK_CUFFT(expr)                                                              \
  do {                                                                                     \
    cufftResult _r = (expr);                                                               \
    if (_r != CUFFT_SUCCESS) {                                                             \
      ::cip::set_error("%s failed: cufft error %d (%s:%d)", #expr, (int)_r, __FILE__, __LINE__); \
      return _r == CUFFT_ALLOC_FAILED ? CIP_ENOMEM : CIP_ECUFFT;                           \
    }                                                                                      \
  } while (0)

#define CIP_TRY(expr)                  \
  do {                                 \
    cip_status _s = (expr);            \
    if (_s != CIP_OK) return _s;       \
  } while (0)

#define CIP_REQUIRE(cond, ...)         \
  do {                                 \
    if (!(cond)) {                     \
      ::cip::set_error(__VA_ARGS__);   \
      return CIP_EINVAL;               \
    }                                  \
  } while (0)

// NVTX range around a stage (visible in Nsight timelines; a no-op when no tool is attached).
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

// Makes the context's device current for the duration of an entry point and restores the caller's device.
struct DeviceScope {
  int prev = -1;
  bool switched = false;
  cudaError_t err = cudaSuccess;
  explicit DeviceScope(int dev) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != dev) {
      err = cudaSetDevice(dev);
      switched = err == cudaSuccess;
    }
  }
  ~DeviceScope() {
    if (switched) cudaSetDevice(prev);
  }
  DeviceScope(const DeviceScope&) = delete;
  DeviceScope& operator=(const DeviceScope&) = delete;
};
#define CIP_ENTER_DEVICE(ctx)                    \
  ::cip::DeviceScope _dev_scope((ctx)->device);  \
  CIP_CHECK_CUDA(_dev_scope.err)

// Grow-only device buffer owned by the context.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cip_status ensure(size_t bytes);
  void release();
  template <typename T> T* as() const { return static_cast<T*>(p); }
};

// Counters of one binning pass, in device memory (and mirrored into mapped host memory by the scan): nothing of
// the binning is read by the host before the end of the call.
struct BinHeader {
  uint32_t ngroups;    // k-groups of 8 records (padded) written by this pass
  uint32_t njobs;      // gridding jobs
  uint32_t overflow;   // != 0: a capacity (jobs, records) was too small -- everything downstream skips its work
  uint32_t pad;
};

// Workspaces of one binning pass (and the records / jobs the gridder reads from it).
struct BinWork {
  DevBuf table;          // u32 per sort key: sample count, then the running record cursor of the key's group
  DevBuf tile_sums;      // per scan tile: (k-groups, jobs), then their exclusive prefixes
  DevBuf jobs, recs, finv, hdr, win_job_start;
  DevBuf items;          // first job of every work item (u32, cap_jobs entries), then the item count and a work counter
  int64_t cap_jobs = 0, cap_groups = 0;  // what jobs / recs currently hold
  void release() {
    for (DevBuf* b : {&table, &tile_sums, &jobs, &recs, &finv, &hdr, &win_job_start, &items}) b->release();
  }
};

// Workspaces of the uvw_tiling binning (tiling.cu).
struct TileWork {
  DevBuf row_nruns, row_off, runs, run_keys, run_keys_alt, run_idx, run_idx_alt, run_len, flags, group_start, cub_tmp;
  void release() {
    for (DevBuf* b : {&row_nruns, &row_off, &runs, &run_keys, &run_keys_alt, &run_idx, &run_idx_alt, &run_len, &flags,
                      &group_start, &cub_tmp})
      b->release();
  }
};

struct FftPlan {
  cufftHandle handle = 0;
  size_t work = 0;
};
struct PfftPlan;  // tables of the own transform (pfft.cu)

// One gridder-ready visibility: exactly one 32-byte sector, so that the scattered writes of the
// record pass never touch a partial sector and the gridder's producer reads two 128-bit words.
//   cell  bits 0-4 i0 (u cell of the support origin inside the tile, 0..kTile-1), 5-9 j0,
//         10-13 first plane relative to the window start, 16-20 / 21-25 origin of the 8 x 8 register
//         window of the gridder (i0, j0 snapped to the lattice of 9-W cells)
//   fu, fv, fw  distance of the first cell / plane from the left edge of the support, in [0,1)
//               (float32 in Rec<float>, 0.32 fixed point in Rec<double>)
//   re, im      weighted, conjugated-if-flipped, n-shift-rotated visibility (R = float or double)
template <typename R>
struct __align__(32) Rec;
template <>
struct __align__(32) Rec<float> {
  uint32_t cell;
  float fu, fv, fw;
  float re, im;
  float pad[2];
};
template <>
struct __align__(32) Rec<double> {
  uint32_t cell;
  uint32_t fu, fv, fw;  // the same fractions as 0.32 fixed point (2.3e-10 of a cell): float32 would cap the
  double re, im;        // accuracy of the float64 path near 1e-7
};
// Packed record of the float32 register-window gridder (W <= 8): 16 bytes, half the scattered write traffic of
// the record pass and half the gridder's record reads.
//   geo   bits 0-2 i0, 3-5 j0 (support origin inside the tile), 6-9 first plane inside the window,
//         10-27 / 28-45 / 46-63 fu, fv, fw as 0.18 fixed point (rounded; 1.9e-6 of a cell, ~1e-5 relative on a
//         kernel weight -- below the float32 path's own rounding)
struct __align__(16) RecP {
  uint64_t geo;
  float re, im;
};
static_assert(sizeof(RecP) == 16, "packed records are half a sector");
constexpr double kFix18 = 262144.0;
constexpr double kFix32 = 4294967296.0;
static_assert(sizeof(Rec<float>) == 32 && sizeof(Rec<double>) == 32, "records are one sector");
inline size_t recs_bytes(int64_t nrec) { return (size_t)nrec * 32; }

// One gridding job: the records of one sort key = one (w window, uv tile, register-window lattice position), i.e.
// one accumulator window of the gridder, padded with zero records to whole k-groups; long ones are cut.
struct __align__(16) Job {
  uint32_t group_begin;  // first k-group (records group_begin * 8 ...)
  uint32_t ngroups;      // <= kJobGroups; bit 31 (kJobStartsItem) marks the first job of a work item
  uint16_t tu, tv;       // uv tile
  uint16_t win;          // window: planes [win * win_step, win * win_step + kWin)
  uint8_t lu, lv;        // lattice position of the register window inside the tile
};
static_assert(sizeof(Job) == 16, "jobs are read with one 128-bit load");

// Geometry shared by the scan / binning / record / gridding kernels.
struct GridGeom {
  double px, py;          // pixel sizes (direction cosines)
  double w0, inv_dw;      // plane p sits at w0 + p*dw
  double nshift;          // n-1 shift s
  int nu, nv;             // oversampled grid
  int W;                  // kernel support
  int nplanes;
  int win_step;           // first-plane stride between windows = kWin - W + 1
  int nwin;               // number of windows
  int ntu, ntv;           // tiles per axis
  int nlat;               // register-window lattice positions per tile edge (1 for wide supports)
  int win_begin, win_end; // windows binned by the current pass (keys of other windows are skipped)
  int do_wstacking;
  float beta_w;           // beta * W
  int bf16_operands;      // float32 tensor path: the kernel's own error (>= 5e-6) allows two-term bfloat16 operands
  double beta_wd;         // the same in float64 (float64 path)
};

}  // namespace cip

struct cip_ctx {
  int device = 0;
  int sm_count = 148;
  size_t smem_optin = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;          // host -> device visibility chunks, overlapped with compute
  cudaEvent_t chunk_ev[cip::kMaxChunks] = {};  // chunk c of the visibilities has arrived
  cudaEvent_t small_ev = nullptr;              // uvw, freq, weights and mask have arrived
  cudaEvent_t sync_ev = nullptr;               // non-timing event for host-clock stage timing
  double t_stage_begin = 0;                    // host clock (ms) when the staging of the current call began
  cudaEvent_t ev[16] = {};
  // hand-over of staged host inputs from cip_scan to the cip_grid_accumulate that follows it
  bool staged_valid = false;
  const void* staged_key[4] = {};
  int64_t staged_dims[4] = {};
  int64_t launches = 0;
  // inputs staged on the device when the caller passes host memory
  cip::DevBuf in_uvw, in_freq, in_vis, in_wgt, in_mask, in_rows;
  cip::DevBuf out_img;
  // scan
  cip::DevBuf scan_out;
  void* scan_host = nullptr;      // pinned + mapped (kSmallHostBytes): small results land here without a copy engine
  void* scan_host_dev = nullptr;  // its device-side address
  // binning workspaces: two sets, so that chunk c+1 is binned (context stream) while chunk c is gridded
  // (grid stream)
  cip::BinWork work[2];
  cip::DevBuf call_flags;         // u32[4]: [0] sticky overflow of the binning passes of the current call; bytes 32..47: u extent
  int rows_kept_blocks = 0;       // > 0: only the first and the last rows_kept_blocks row blocks of every resident plane hold
                                  // data (the others were not cleared and must not be read); 0: all rows
  void* flags_host = nullptr;     // mapped pinned mirror of call_flags and the last BinHeader (read after the sync)
  void* flags_host_dev = nullptr;
  cudaStream_t grid_stream = nullptr;
  cudaEvent_t binned_ev[2] = {}, gridded_ev[2] = {}, zero_ev = nullptr;  // non-timing
  cudaEvent_t chunk_t[cip::kMaxChunks][4] = {};                          // timing: bin begin/end, grid begin/end
  // grids, fft, image
  cip::DevBuf grid, image, tables, fft_t;  // fft_t: one transposed, cropped plane (ny x nu)
  std::vector<double> tables_key;  // plan parameters the correction tables were built for
  int tables_ncheb = 0;
  std::map<std::tuple<int, int, int, int>, cip::FftPlan> fft_plans;  // (nu, nv, batch, fp64)
  std::map<std::tuple<int, int, int>, cip::PfftPlan*> pfft_plans;     // (nu, nv, fp64)
  // tiling
  cip::DevBuf t_uvw, t_freq, t_vis;
  cip::TileWork twork;
};

namespace cip {

// ---- planning (plan.cpp) ----------------------------------------------------------
int good_size(int n);
struct KernelEntry { double beta, err; };
bool kernel_lookup(double sigma, int W, KernelEntry* out);
void gauss_legendre(int n, std::vector<double>& x, std::vector<double>& w);
// c(xi) = int_{-W/2}^{W/2} phi(2t/W) cos(2 pi t xi) dt
double kernel_ft_one(double xi, int W, double beta, const std::vector<double>& glx, const std::vector<double>& glw);
int correction_quadrature_size(int W);
GridGeom make_geom(const cip_plan& p);

// ---- staged inputs (ctx.cu) -------------------------------------------------------
struct DevInputs {
  const double* uvw;
  const double* freq;
  const float2* vis;   // wide: really const double2*
  const float* wgt;    // wide: really const double*
  const uint8_t* mask;
  int wide = 0;        // 1: complex128 visibilities and float64 weights (cip_invert_args.vis_c128)
  int64_t nrow, nchan, vstride, wstride;
  double w_lo = 0.0, w_hi = 0.0;  // w_hi > 0: only samples with w_lo <= |w| < w_hi are active
  // jagged rows (nullptr: dense): channels [ch0[r], ch1[r]) of row r at elements voff[r] ... of vis / wgt / mask
  const int64_t* voff = nullptr;
  const int32_t* ch0 = nullptr;
  const int32_t* ch1 = nullptr;
  int64_t nvis = 0;
};
cip_status validate_args(const cip_invert_args* a);
// Rows [row0[c], row0[c+1]) of the visibilities are on the device once ready[c] has fired.
struct VisFeed {
  int nchunks = 0;
  int64_t row0[kMaxChunks + 1] = {};
  cudaEvent_t ready[kMaxChunks] = {};
};
// Host inputs (memspace 0): uvw, freq, weights and mask are copied on the context stream (skipped when
// `reuse_small` and the previous call was a cip_scan of the same arrays); when `feed` is given the
// visibilities follow in row chunks on the copy stream, else they are not copied at all (cip_scan).
// Device inputs (memspace 1) are used in place and the feed, if any, is a single chunk that is ready.
cip_status stage_inputs(cip_ctx* ctx, const cip_invert_args* a, DevInputs* d, VisFeed* feed, bool reuse_small,
                        double* ms_h2d);
// Make the context stream wait for every chunk of the feed (for code that needs all visibilities).
cip_status wait_feed(cip_ctx* ctx, const VisFeed& feed);
float elapsed_ms(cip_ctx* ctx, int e0, int e1);
// Bring `bytes` (a multiple of 4, <= kSmallHostBytes) of device memory to the host and wait for it.  A tiny
// kernel writes them into mapped pinned memory: a cudaMemcpy here would queue on a copy engine
// behind the visibility chunks that are streaming in and stall the compute stream until they are done.
constexpr size_t kSmallHostBytes = 64 << 10;
cip_status fetch_small(cip_ctx* ctx, const void* dev, size_t bytes, void* host_out);

// ---- stages -----------------------------------------------------------------------
cip_status run_scan(cip_ctx* ctx, const DevInputs& in, int do_wstacking, cip_scan_result* out);
cip_status run_w_histogram(cip_ctx* ctx, const DevInputs& in, double wmin, double wmax, int nbins, int64_t* counts_out);

// Bin the samples whose window lies in [g.win_begin, g.win_end) into workspace `slot`: records + jobs + header,
// all on the device (see bin.cu).  `call_flags` (device, u32) collects overflow over all passes of a call.
cip_status run_binning(cip_ctx* ctx, int slot, const DevInputs& in, const GridGeom& g, int fp64, bool want_items);
// Which float32 W <= 8 gridder runs: true = the tcgen05 kernel (grid_tc.cu, jobs), false = the register-window mma.sync
// kernel (grid.cu, work items).
bool use_tc_gridder();
// Largest number of windows one binning pass may cover (bounded by the size of the key table).
int max_windows_per_pass(const GridGeom& g);
// After the stream has drained: did any binning pass of the call overflow its buffers?  Grows them if so.
cip_status binning_overflowed(cip_ctx* ctx, bool* again);

// Grid the jobs of windows [win_first, win_last] of binning workspace `slot` into `grid` (planes slot = plane -
// plane_base), on `stream`.
cip_status run_gridder(cip_ctx* ctx, int slot, cudaStream_t stream, const GridGeom& g, int fp64, int win_first,
                       int win_last, void* grid, int plane_base, int nslots);
cip_status run_fft(cip_ctx* ctx, void* grid, int nu, int nv, int batch, int fp64, int keep_v = 0);
cip_status fft_rows(cip_ctx* ctx, void* data, int n, int batch, int fp64);  // data == nullptr: plan only
// FFT + w-screen of the resident planes [plane_base, plane_base + nslots) into the image.  Default: the own pruned
// transform with the fused w-screen (pfft.cu) where the row fits shared memory.  Else, and with CIP_FFT=cufft: per
// plane, cuFFT 1-D transforms along v, transpose of the retained columns, 1-D transforms along u, w-screen from the
// transposed plane (see image.cu); CIP_FFT_2D=1: cuFFT's 2-D plan on every plane, then one w-screen pass.
cip_status run_fft_screen(cip_ctx* ctx, const cip_plan& p, void* grid, int plane_base, int nslots, void* image,
                          double* ms_fft, double* ms_wscreen);
cip_status pfft_prepare(cip_ctx* ctx, const cip_plan& p, bool* ok);  // tables and work area; *ok: the plan is supported
// Row blocks (pfft.cu): a plane of 128 MB and more counts as kRowBlocks blocks of rows, and the blocks the visibilities
// cannot reach (|u| beyond their extent: rows that were never written) are neither cleared, nor transformed along v,
// nor read.  pfft_row_blocks: rows per block if this plan's transform can skip blocks, else 0.
constexpr int kRowBlocks = 32;
int pfft_row_blocks(cip_ctx* ctx, const cip_plan& p);
cip_status run_pfft_screen(cip_ctx* ctx, const cip_plan& p, void* grid, int plane_base, int nslots, void* image);
void pfft_release(cip_ctx* ctx);
cip_status run_wscreen(cip_ctx* ctx, const cip_plan& p, const void* grid, int plane_base, int nslots, void* image);
cip_status run_correct(cip_ctx* ctx, const cip_plan& p, void* image, int row_begin, int row_end, double scale);

}  // namespace cip
