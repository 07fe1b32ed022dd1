/* cip.h -- C-ABI of the B200-native invert hot path (libcip_b200.so).
 *
 * Drop-in boundary for the step the reference delegates to native code:
 *
 *   reference interface                                           replaced by
 *   ------------------------------------------------------------  -------------------------
 *   ducc0.wgridder.ms2dirty(uvw, freq, ms, wgt, npix_x, npix_y,    cip_invert
 *     pixsize_x, pixsize_y, epsilon=, do_wstacking=, nthreads=,    (cip_scan + cip_plan_make +
 *     mask=)   call site src/ska_sdp_cip/invert.py:170-183          cip_grid_accumulate +
 *                                                                   cip_correct_image when the
 *                                                                   visibilities are sharded)
 *   effective_weights.sum()         src/ska_sdp_cip/invert.py:184  cip_scan_result.sum_wgt
 *   StokesIGridderInput.from_measurement_set_reader / effective_weights
 *                                   src/ska_sdp_cip/invert.py:72-116   cip_stokes_i
 *   create_uvw_tile_mapping(uvw, tile_size, channel_freqs)         cip_tile_bin
 *     src/ska_sdp_cip/uvw_tiling/tiling_plan.py:84-134 (and :29-61, :150-181)
 *   Tile._from_jagged_visibilities_slice                           cip_tile_gather
 *     src/ska_sdp_cip/uvw_tiling/tile.py:83-115
 *   integrate_weighted_images (sum of chunk images / sum of weights)
 *     src/ska_sdp_cip/invert.py:200-209                            caller-side NCCL reduce of the
 *                                                                   uncorrected images, then
 *                                                                   cip_correct_image(scale=1/sum)
 *
 * Conventions: plain pointers and sizes only; the caller owns every input and
 * output buffer; the library owns what hangs off cip_ctx (streams, cuFFT plans,
 * device workspaces).  A cip_ctx is bound to one CUDA device and is not
 * re-entrant (one context per process and GPU).  Every entry point returns a
 * cip_status; cip_last_error() gives a thread-local message for the last
 * failure.  There is no CPU fallback: without a CUDA device cip_ctx_create fails.
 */
#ifndef CIP_H
#define CIP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cip_ctx cip_ctx;

typedef enum {
  CIP_OK = 0,
  CIP_EINVAL = 1, /* bad argument (Python shim raises ValueError)            */
  CIP_ENOMEM = 2, /* device memory exhausted                                 */
  CIP_ECUDA = 3,  /* CUDA runtime error, incl. "no device" (RuntimeError)    */
  CIP_ECUFFT = 4  /* cuFFT error                                             */
} cip_status;

const char* cip_last_error(void);
const char* cip_version(void);

/* device < 0: use the current CUDA device */
cip_status cip_ctx_create(int device, cip_ctx** out);
cip_status cip_ctx_destroy(cip_ctx* ctx);
/* The stream all work of this context is enqueued on (a cudaStream_t). */
void* cip_ctx_stream(cip_ctx* ctx);
cip_status cip_ctx_synchronize(cip_ctx* ctx);
/* Number of kernels this library has launched on the context so far (own kernels, CUB passes, cuFFT execs). */
int64_t cip_ctx_launches(cip_ctx* ctx);
/* The CUDA device index the context is bound to.  Every entry point makes it current for the call and restores
 * the caller's current device before it returns. */
int cip_ctx_device(cip_ctx* ctx);
/* The calling thread's current CUDA device (what device < 0 resolves to). */
cip_status cip_current_device(int* device);

/* ---- ms2dirty arguments, 1:1 (invert.py:170-183) -------------------------------- */
typedef struct {
  const double* uvw;        /* (nrow,3) C order, metres                              */
  int64_t nrow;
  const double* freq;       /* (nchan,) Hz                                           */
  int64_t nchan;
  const void* vis;          /* (nrow,nchan) complex64 (complex128 if vis_c128)       */
  const void* wgt;          /* (nrow,nchan) float32 (float64 if vis_c128) or NULL (= all ones) */
  const uint8_t* mask;      /* (nrow,nchan) or NULL; 0 = skip sample                 */
  int64_t vis_row_stride;   /* elements between rows of vis (>= nchan; 0 = nchan)    */
  int64_t wgt_row_stride;   /* same for wgt and mask                                 */
  int npix_x, npix_y;       /* even, >= 16                                           */
  double pixsize_x, pixsize_y; /* radians (direction cosines per pixel)              */
  double epsilon;
  int do_wstacking;         /* 1 = improved w-stacking, 0 = ignore w                 */
  int accumulate_fp64;      /* 0: complex64 grids; 1: complex128 grids + tighter kernel */
  int memspace;             /* 0: all pointers are host memory; 1: device memory of the ctx device */
  /* Jagged rows (all NULL / 0 for the dense (nrow,nchan) layout): row r holds the channels
   * [row_chan_start[r], row_chan_stop[r]) only, its visibilities (and weights, mask) are the row_chan_stop[r] -
   * row_chan_start[r] elements starting at element row_vis_offset[r] of vis (wgt, mask), nvis elements in total.
   * This is the layout of the reference's uvw tiles (Tile.uvw / visibilities / channel_start_indices /
   * channel_stop_indices, src/ska_sdp_cip/uvw_tiling/tile.py:14-50): tiles are gridded as they are stored.   */
  const int64_t* row_vis_offset;
  const int32_t* row_chan_start;
  const int32_t* row_chan_stop;
  int64_t nvis;
  double w_lo, w_hi;        /* w_hi > 0: only the samples with w_lo <= |w| < w_hi (wavelengths) take part -- one
                             * |w| slab of a w-sharded invert (SURVEY 8e option 3); w_hi <= 0: all samples      */
  int32_t vis_c128;         /* 1: vis is complex128 and wgt (if given) float64 -- ducc0's double-precision call.  The
                             * samples are read as they are (no down-cast anywhere: the records of the float64
                             * gridder carry float64 visibilities); combine with accumulate_fp64 = 1 for ducc0's
                             * behaviour on complex128 input.  0: complex64 / float32                              */
} cip_invert_args;

typedef struct {
  double wmin, wmax;        /* range of |w| in wavelengths over active samples       */
  int64_t nactive;          /* samples with wgt*mask != 0 (the visibilities are not read) */
  double sum_wgt;           /* sum of wgt*mask over ALL samples (invert.py:184)      */
} cip_scan_result;

typedef struct {
  int npix_x, npix_y;
  double pixsize_x, pixsize_y;
  double sigma;             /* oversampling actually used                            */
  int support;              /* kernel support W in cells                             */
  double beta;              /* phi(x) = exp(beta*W*(sqrt(1-x^2)-1))                   */
  int nu, nv;               /* oversampled grid                                      */
  double nshift;            /* n-1 shift s                                           */
  double dw, w0;            /* plane p sits at w0 + p*dw                              */
  int nplanes;
  int accumulate_fp64;
  int do_wstacking;
  int plane_block;          /* planes resident at once (filled by the library)       */
} cip_plan;

typedef struct {
  double ms_total, ms_h2d, ms_scan, ms_bin, ms_grid /* gridding kernel only */, ms_fft, ms_wscreen, ms_correct, ms_d2h;
  int64_t launches;         /* kernels of this library launched by the call          */
  int64_t nactive;
  int nblocks;              /* plane blocks processed = launches of the gridding kernel */
  double ms_zero;           /* clearing the plane stack before gridding                 */
} cip_report;

/* Host-only planning: choose (sigma, W) and the w-plane layout.  No GPU needed.
 * sigma_override/support_override: 0 = automatic. */
cip_status cip_plan_make(int npix_x, int npix_y, double pixsize_x, double pixsize_y, double epsilon,
                         int do_wstacking, int accumulate_fp64, double wmin, double wmax, int64_t nactive,
                         double sigma_override, int support_override, cip_plan* out);

/* Kernel design table of the library: rms aliasing error and beta for (sigma, W). */
cip_status cip_kernel_table(double sigma, int support, double* beta, double* rms_err);

/* Gridding-kernel Fourier transform c(xi) used by the correction (host, float64). */
cip_status cip_kernel_ft(const cip_plan* plan, const double* xi, int64_t n, double* out);

/* G1 scan: w-range, active count and weight sum of one shard.  With host arguments (memspace 0) the staged
 * uvw / freq / weights / mask are reused by a cip_grid_accumulate that follows with the SAME pointers and
 * sizes: the arrays must not change in between. */
cip_status cip_scan(cip_ctx* ctx, const cip_invert_args* args, cip_scan_result* out);

/* |w| histogram of the active samples of one shard: nbins equal bins over [wmin, wmax], counts ADDED to counts_out
 * (host, int64[nbins]).  The ranks of a w-sharded invert sum their histograms and cut the |w| axis into slabs of
 * equal population; each rank then grids one slab (args->w_lo, w_hi) with a plan of its own w range. */
cip_status cip_w_histogram(cip_ctx* ctx, const cip_invert_args* args, double wmin, double wmax, int nbins,
                           int64_t* counts_out);

/* The same histogram for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Device inputs only (args->memspace = 1, else CIP_EINVAL); counts_out is host memory, int64[nbins], counts
 * ADDED.  The kernel's per-bin counters live in a static device array of CIP_W_HISTOGRAM_STATIC_MAX_BINS x 8 bytes
 * (32 KB held by every GPU that loads the library, until exit): nbins beyond that is rejected with CIP_EINVAL.  The call
 * makes no cudaMalloc / cudaFree and calls no allocating library routine (the first call on a device resolves the
 * array's address).  All calls on a device share the array -- each call clears the bins it uses, and calls that overlap
 * on the GPU (two contexts / streams of the same device at once) race. */
#define CIP_W_HISTOGRAM_STATIC_MAX_BINS 4096
cip_status cip_w_histogram_static(cip_ctx* ctx, const cip_invert_args* args, double wmin, double wmax, int nbins,
                                  int64_t* counts_out);

/* G1 binning + G2 gridding + G3 FFT + G4 w-screen: ADD the uncorrected dirty image of the
 * samples in args to `image` (device memory, (npix_x,npix_y) float32, or float64 when
 * plan->accumulate_fp64).  Linear in the samples, so shards may be reduced before
 * cip_correct_image. */
cip_status cip_grid_accumulate(cip_ctx* ctx, const cip_invert_args* args, cip_plan* plan, void* image_dev,
                               cip_report* report /* may be NULL */);

/* Inspection of stage G1 (tests; not needed by callers): bin the samples of `args` for `plan` and copy the gridding
 * jobs and the gridder-ready records to host memory.  Call with jobs_out = recs_out = NULL for the counts first.
 * A job is the group of records of one (w window, uv tile, register-window lattice position), padded with zero
 * records to a multiple of 8; record r of job j is records[(jobs[j].group_begin * 8 + r) * rec_bytes].
 * rec_bytes = 16: { u64 geo: bits 0-2 i0, 3-5 j0 (support origin inside the tile), 6-9 first plane inside the window,
 *                   10-27 / 28-45 / 46-63 fractional offsets fu, fv, fw as 0.18 fixed point; f32 re, im }
 * rec_bytes = 32: { u32 cell: bits 0-4 i0, 5-9 j0, 10+ first plane; fu, fv, fw as f32 (float32 grids) or 0.32 fixed point
 *                   (float64 grids); re, im as f32 + 8 bytes pad, or f64 } */
typedef struct {
  uint32_t group_begin, ngroups; /* ngroups: bit 31 is a flag, mask it */
  uint16_t tu, tv, win;
  uint8_t lu, lv;
} cip_job;
cip_status cip_bin_inspect(cip_ctx* ctx, const cip_invert_args* args, const cip_plan* plan, int64_t* njobs,
                           int64_t* ngroups, int* rec_bytes, cip_job* jobs_out, void* recs_out);

/* Stage G3 on its own (tests; callers that want the transformed plane): the nx x ny cells of the unnormalised
 * e^{+i} 2-D transform of one nu x nv plane (complex64, or complex128 if fp64) that the image keeps,
 *   out[i][j] = sum_{u,v} plane[u][v] exp(+2 pi i (u fu / nu + v fv / nv)),  fu = (i - nx/2) mod nu, fv = (j - ny/2) mod nv,
 * computed by the pruned transform of csrc/pfft.cu (what ducc0 does inside ms2dirty between gridding and the
 * w-screen; reference call site src/ska_sdp_cip/invert.py:170-183).  plane and out are device pointers; the plane is
 * overwritten (it holds the row transforms afterwards). */
cip_status cip_plane_fft_crop(cip_ctx* ctx, const void* plane, int nu, int nv, int nx, int ny, int fp64, void* out);

/* The same for several device-resident shards (memspace 1; e.g. the chunks of the reference's partition that one
 * GPU owns, or what a |w| slab received from the other ranks): all shards are gridded into ONE plane stack, which
 * is transformed once -- instead of one set of per-plane FFTs per shard. */
cip_status cip_grid_accumulate_many(cip_ctx* ctx, const cip_invert_args* args, int nshards, cip_plan* plan,
                                    void* image_dev, cip_report* report /* may be NULL */);

/* G4 final + P4: image[r,:] *= scale / (c_u c_v c_w n) for rows row_begin..row_end-1.
 * `image_dev` points at row 0 of the full image. */
cip_status cip_correct_image(cip_ctx* ctx, const cip_plan* plan, void* image_dev, int row_begin, int row_end,
                             double scale);

/* The whole of ms2dirty on one GPU.  dirty_out: (npix_x,npix_y) float32 (float64 when
 * accumulate_fp64) in the memory space of args->memspace.  scan_out, plan_out, report may be NULL. */
cip_status cip_invert(cip_ctx* ctx, const cip_invert_args* args, void* dirty_out, cip_scan_result* scan_out,
                      cip_plan* plan_out, cip_report* report);

/* ---- P1/P2: Stokes-I prologue (invert.py:72-116) -------------------------------- */
/* vis4 (nrow,nchan,4) complex64, flags4 (nrow,nchan,4) uint8/bool, wgt4 (nrow,nchan,4) float32
 * -> vis_I = 0.5 (v0+v3), eff_w = (!(f0|f3)) * 4/(1/w0+1/w3).  memspace as above. */
cip_status cip_stokes_i(cip_ctx* ctx, const void* vis4, const uint8_t* flags4, const float* wgt4,
                        int64_t nrow, int64_t nchan, void* vis_out, float* effw_out, int memspace);

/* The same on device memory, enqueued on cip_ctx_stream without waiting: for callers that pipeline their own
 * uploads (order them against the stream with events). */
cip_status cip_stokes_i_async(cip_ctx* ctx, const void* vis4, const uint8_t* flags4, const float* wgt4, int64_t nrow,
                              int64_t nchan, void* vis_out, float* effw_out);

/* ---- T1-T3: uvw tile binning (tiling_plan.py:29-181), bit-exact ------------------ */
typedef struct cip_tile_result cip_tile_result; /* device-resident result of one binning */

/* idx = (int64) floor((freq/c) * (uvw/tile_size) + 0.5), run-length compressed per row,
 * grouped by tile in first-appearance order, slices ascending (irow, chan_start). */
cip_status cip_tile_bin(cip_ctx* ctx, const double* uvw, int64_t nrow, const double* freq, int64_t nchan,
                        const double tile_size[3], int64_t row_offset, int memspace,
                        cip_tile_result** result, int64_t* n_tiles, int64_t* n_slices);
/* Copy the result out (host buffers; any pointer may be NULL). */
cip_status cip_tile_fetch(cip_ctx* ctx, const cip_tile_result* result,
                          int64_t* tile_coords /* (n_tiles,3) */, int64_t* tile_offsets /* (n_tiles+1,) */,
                          int64_t* slice_irow, int32_t* slice_ch_start, int32_t* slice_ch_stop /* (n_slices,) */);
/* T5 gather (tile.py:83-115): visibilities (nrow,nchan) complex64 reordered tile-contiguous.
 * vis_out (nvis_total,) complex64, vis_offsets (n_slices+1,) int64, uvw_out (n_slices,3) float64. */
cip_status cip_tile_gather(cip_ctx* ctx, const cip_tile_result* result, const void* vis, const double* uvw,
                           int64_t nrow, int64_t nchan, int memspace, void* vis_out, int64_t* vis_offsets,
                           double* uvw_out);
cip_status cip_tile_free(cip_ctx* ctx, cip_tile_result* result);

#ifdef __cplusplus
}
#endif
#endif /* CIP_H */
