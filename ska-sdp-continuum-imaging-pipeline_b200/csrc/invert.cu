// The ms2dirty replacement: scan -> plan -> binning -> per block of w planes { grid, FFT,
// w-screen } -> correction.  Replaces the single native call of the reference,
// ducc0.wgridder.ms2dirty at src/ska_sdp_cip/invert.py:170-183, and the weight sum of :184.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.h"

using namespace cip;

namespace {

// Stage timing for cip_report.  Device inputs: CUDA timing events on the context stream.  Host inputs:
// while visibility chunks are in flight on the copy stream NO timing event may be recorded on the
// compute stream -- a timing-enabled cudaEventRecord is ordered behind all copy-engine work already
// submitted on other streams (measured on B200, driver 580: tools/experiments/overlap_probe.cu mode 9),
// which would serialise the whole compute stream behind the upload.  So host inputs are timed with
// the host clock around a non-timing event synchronisation (every stage ends with one anyway).
struct StageTimer {
  cip_ctx* ctx;
  bool on, host;
  double t0 = 0;
  StageTimer(cip_ctx* c, bool enabled, bool host_clock) : ctx(c), on(enabled), host(host_clock) {}
  void start() {
    if (!on) return;
    if (host) t0 = now_ms();
    else cudaEventRecord(ctx->ev[0], ctx->stream);
  }
  void stop(double* acc) {
    if (!on) return;
    if (host) {
      cudaEventRecord(ctx->sync_ev, ctx->stream);
      cudaEventSynchronize(ctx->sync_ev);
      *acc += now_ms() - t0;
    } else {
      cudaEventRecord(ctx->ev[1], ctx->stream);
      cudaEventSynchronize(ctx->ev[1]);
      *acc += elapsed_ms(ctx, 0, 1);
    }
  }
};

// Host -> device copies of the caller's arrays may still be in flight on the copy stream when an entry
// point returns early (error, nothing active): never return before they have drained.
struct CopyStreamGuard {
  cip_ctx* ctx;
  explicit CopyStreamGuard(cip_ctx* c) : ctx(c) {}
  ~CopyStreamGuard() { cudaStreamSynchronize(ctx->copy_stream); }
};



}  // namespace

extern "C" cip_status cip_scan(cip_ctx* ctx, const cip_invert_args* args, cip_scan_result* out) {
  CIP_REQUIRE(ctx && out, "NULL argument");
  CIP_TRY(validate_args(args));
  CIP_ENTER_DEVICE(ctx);
  DevInputs in;
  CIP_TRY(stage_inputs(ctx, args, &in, nullptr, false, nullptr));  // no visibilities: the scan reads weights only
  CIP_TRY(run_scan(ctx, in, args->do_wstacking, out));
  ctx->staged_valid = args->memspace == 0;  // a cip_grid_accumulate of the same arrays may reuse them
  return CIP_OK;
}

extern "C" cip_status cip_w_histogram(cip_ctx* ctx, const cip_invert_args* args, double wmin, double wmax, int nbins,
                                      int64_t* counts_out) {
  CIP_REQUIRE(ctx && counts_out, "NULL argument");
  CIP_TRY(validate_args(args));
  CIP_ENTER_DEVICE(ctx);
  DevInputs in;
  CIP_TRY(stage_inputs(ctx, args, &in, nullptr, true, nullptr));  // weights and uvw only, as for the scan
  CIP_TRY(run_w_histogram(ctx, in, wmin, wmax, nbins, counts_out));
  ctx->staged_valid = args->memspace == 0;
  return CIP_OK;
}

// Largest |u| (metres) over the rows and largest frequency, as the bit patterns of non-negative doubles (they order like
// unsigned integers): what bounds the plane rows the visibilities can reach.
__global__ void __launch_bounds__(256) uext_kernel(const double* __restrict__ uvw, int64_t nrow, const double* __restrict__ freq,
                                                   int64_t nchan, unsigned long long* __restrict__ out) {
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  double um = 0.0, fm = 0.0;
  for (int64_t r = t0; r < nrow; r += stride) um = fmax(um, fabs(uvw[3 * r]));
  for (int64_t c = t0; c < nchan; c += stride) fm = fmax(fm, fabs(freq[c]));
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    um = fmax(um, __shfl_xor_sync(0xffffffffu, um, o));
    fm = fmax(fm, __shfl_xor_sync(0xffffffffu, fm, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (um > 0.0) atomicMax(out, (unsigned long long)__double_as_longlong(um));
    if (fm > 0.0) atomicMax(out + 1, (unsigned long long)__double_as_longlong(fm));
  }
}

// How many row blocks at either end of a plane (FFT order: u = 0 is row 0) the shards' visibilities can reach: a sample
// at u wavelengths touches the rows within W / 2 + 1 of u px nu.  0: no block can be skipped (or the transform does
// not work block-wise).  One small kernel and one 16-byte read-back per call.
static cip_status reachable_row_blocks(cip_ctx* ctx, const DevInputs* ins, int nshards, const cip_plan& plan, int* kept) {
  *kept = 0;
  const int bs = pfft_row_blocks(ctx, plan);
  if (bs == 0) return CIP_OK;
  unsigned long long* d = reinterpret_cast<unsigned long long*>(static_cast<char*>(ctx->call_flags.p) + 32);
  CIP_CHECK_CUDA(cudaMemsetAsync(d, 0, 16, ctx->stream));
  for (int k = 0; k < nshards; ++k) {
    if (ins[k].nrow == 0) continue;
    const int blocks = (int)std::min<int64_t>((ins[k].nrow + 255) / 256, 4 * ctx->sm_count);
    uext_kernel<<<blocks, 256, 0, ctx->stream>>>(ins[k].uvw, ins[k].nrow, ins[k].freq, ins[k].nchan, d);
    ctx->launches++;
  }
  CIP_CHECK_CUDA(cudaGetLastError());
  double ext[2];
  CIP_TRY(fetch_small(ctx, d, 16, ext));
  const double cells = ext[0] * ext[1] / kSpeedOfLight * plan.pixsize_x * plan.nu;  // |u| in cells, at most
  if (!(cells < 0.5 * plan.nu)) return CIP_OK;                                       // wraps around (or NaN): every row
  const int rows = (int)std::ceil(cells) + plan.support / 2 + 12;  // support, the 8-cell register window, slack
  const int kb = rows / bs + 1;
  if (2 * kb < kRowBlocks) *kept = kb;
  return CIP_OK;
}

// Rows [r0, r1) of the staged inputs as inputs of their own.
static DevInputs row_slice(const DevInputs& in, int64_t r0, int64_t r1) {
  DevInputs c = in;
  c.nrow = r1 - r0;
  c.uvw = in.uvw + 3 * r0;
  if (in.voff) {  // jagged rows: the offsets are absolute, only the row tables move
    c.voff = in.voff + r0;
    c.ch0 = in.ch0 + r0;
    c.ch1 = in.ch1 + r0;
    return c;
  }
  const int64_t ws = in.wide ? 2 : 1;  // the typed pointers count float2 / float elements
  c.vis = in.vis + ws * r0 * in.vstride;
  if (in.wgt) c.wgt = in.wgt + ws * r0 * in.wstride;
  if (in.mask) c.mask = in.mask + r0 * in.wstride;
  return c;
}

// One attempt: everything is enqueued without a single read-back; `rep` stage times come from event pairs (device
// inputs) or the host clock around event synchronisations (streamed host inputs, see StageTimer).
static cip_status grid_accumulate_attempt(cip_ctx* ctx, const DevInputs* ins, const VisFeed* feeds, int nshards, cip_plan* plan,
                                          void* image_dev, cip_report* rep) {
  const int fp64 = plan->accumulate_fp64;
  GridGeom g = make_geom(*plan);
  const bool want_items = !fp64 && g.W <= kFastSupport && !use_tc_gridder();
  const DevInputs& in = ins[0];
  const VisFeed& feed = feeds[0];
  StageTimer tm(ctx, rep != nullptr, feed.nchunks > 0 && feed.ready[0] != nullptr);
  double dummy = 0;

  // how many planes fit
  CIP_TRY(run_fft_screen(ctx, *plan, nullptr, 0, 0, nullptr, nullptr, nullptr));  // creates and caches the plans
  const size_t plane_bytes = (size_t)plan->nu * plan->nv * (fp64 ? 16 : 8);
  // cudaMemGetInfo is a device-wide driver call (it can wait behind in-flight copies): only ask when the
  // plane stack has to grow
  int64_t max_slots = (int64_t)(ctx->grid.cap / plane_bytes);
  if (max_slots < plan->nplanes) {
    size_t free_b = 0, total_b = 0;
    CIP_CHECK_CUDA(cudaMemGetInfo(&free_b, &total_b));
    // keep room for the binning workspaces (records 16 .. 32 B per visibility, padded) if they are not allocated yet
    size_t reserve = (size_t)768 << 20;
    size_t nvis = 0;
    for (int k = 0; k < nshards; ++k) nvis = std::max(nvis, (size_t)ins[k].nrow * ins[k].nchan);
    if (ctx->work[0].recs.cap < 36 * nvis / std::max(feed.nchunks, 1)) reserve += 40 * nvis;
    size_t avail = free_b + ctx->grid.cap;
    avail = avail > reserve ? avail - reserve : 0;
    max_slots = (int64_t)(avail / plane_bytes);
  }
  if (const char* e = getenv("CIP_MAX_PLANE_SLOTS")) max_slots = std::min<int64_t>(max_slots, atoll(e));
  int nslots_alloc;
  if (max_slots >= plan->nplanes) {
    nslots_alloc = plan->nplanes;
  } else {
    if (max_slots < (g.W <= kFastSupport ? kWin : 1)) {
      set_error("not enough device memory for one window of %d planes of %zu bytes", kWin, plane_bytes);
      return CIP_ENOMEM;
    }
    nslots_alloc = (int)max_slots;
  }
  plan->plane_block = nslots_alloc;
  CIP_TRY(ctx->grid.ensure(plane_bytes * nslots_alloc));
  trace_mark("  grid ensured");
  const int wpp = max_windows_per_pass(g);  // windows one binning pass may cover (size of the key table)

  if (nslots_alloc >= plan->nplanes) {
    // All planes resident: bin and grid each row chunk as soon as its visibilities have arrived (host
    // inputs stream in on the copy stream; device inputs are one chunk), then transform once.
    // Pass k is binned on the context stream into workspace k & 1 and gridded on the grid stream, so that
    // the binning of the next pass overlaps the gridding of this one.  Device inputs are timed with event pairs
    // that nobody waits for until the end; host inputs (timed with the host clock, see StageTimer) give up
    // the overlap when a report is requested.
    const bool dev_timing = rep && !tm.host;
    cudaStream_t gs = ctx->grid_stream;
    // Large planes: the row blocks no visibility reaches are neither cleared, nor transformed along v, nor read
    int kept_blocks = 0;
    CIP_TRY(reachable_row_blocks(ctx, ins, nshards, *plan, &kept_blocks));
    struct KeptGuard {
      cip_ctx* c;
      ~KeptGuard() { c->rows_kept_blocks = 0; }
    } kept_guard{ctx};
    ctx->rows_kept_blocks = kept_blocks;
    // The plane stack is cleared on the grid stream (ordered behind everything already on the context stream):
    // the first gridding pass needs it, the first binning pass does not, and the two have different limits (HBM
    // writes against issue slots), so they run side by side.
    CIP_CHECK_CUDA(cudaEventRecord(ctx->zero_ev, ctx->stream));
    CIP_CHECK_CUDA(cudaStreamWaitEvent(gs, ctx->zero_ev, 0));
    if (dev_timing) CIP_CHECK_CUDA(cudaEventRecord(ctx->ev[14], gs));
    const double tz0 = now_ms();
    if (kept_blocks > 0) {
      // the last kept rows of plane p and the first kept rows of plane p + 1 are one contiguous range
      const size_t edge = plane_bytes / kRowBlocks * kept_blocks;
      char* base = static_cast<char*>(ctx->grid.p);
      CIP_CHECK_CUDA(cudaMemsetAsync(base, 0, edge, gs));
      for (int p = 1; p < plan->nplanes; ++p) CIP_CHECK_CUDA(cudaMemsetAsync(base + plane_bytes * p - edge, 0, 2 * edge, gs));
      CIP_CHECK_CUDA(cudaMemsetAsync(base + plane_bytes * plan->nplanes - edge, 0, edge, gs));
    } else {
      CIP_CHECK_CUDA(cudaMemsetAsync(ctx->grid.p, 0, plane_bytes * plan->nplanes, gs));
    }
    if (dev_timing) CIP_CHECK_CUDA(cudaEventRecord(ctx->ev[15], gs));
    if (rep && !dev_timing) {
      CIP_CHECK_CUDA(cudaEventRecord(ctx->sync_ev, gs));
      CIP_CHECK_CUDA(cudaEventSynchronize(ctx->sync_ev));
      rep->ms_zero += now_ms() - tz0;
    }
    bool used[2] = {false, false};
    int pass = 0, ntimed = 0;
    for (int k = 0; k < nshards; ++k) {
    const DevInputs& in = ins[k];
    const VisFeed& feed = feeds[k];
    for (int c = 0; c < feed.nchunks; ++c) {
      if (feed.row0[c + 1] <= feed.row0[c]) continue;
      if (feed.ready[c]) {
        if (rep) {  // timed call: the host waits for the chunk so that the stage times exclude the upload
          CIP_CHECK_CUDA(cudaEventSynchronize(feed.ready[c]));
          if (c == feed.nchunks - 1) rep->ms_h2d = now_ms() - ctx->t_stage_begin;  // all inputs on the device
        } else {
          CIP_CHECK_CUDA(cudaStreamWaitEvent(ctx->stream, feed.ready[c], 0));
        }
      }
      const DevInputs chunk = row_slice(in, feed.row0[c], feed.row0[c + 1]);
      for (int wa = 0; wa < g.nwin; wa += wpp, ++pass) {
        const int slot = pass & 1;
        g.win_begin = wa;
        g.win_end = std::min(g.nwin, wa + wpp);
        if (used[slot]) CIP_CHECK_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->gridded_ev[slot], 0));  // workspace in use
        const bool timed = dev_timing && ntimed < kMaxChunks;
        if (timed) CIP_CHECK_CUDA(cudaEventRecord(ctx->chunk_t[ntimed][0], ctx->stream));
        else tm.start();
        CIP_TRY(run_binning(ctx, slot, chunk, g, fp64, want_items));
        if (timed) CIP_CHECK_CUDA(cudaEventRecord(ctx->chunk_t[ntimed][1], ctx->stream));
        else tm.stop(rep ? &rep->ms_bin : &dummy);
        trace_mark("  pass binned");
        CIP_CHECK_CUDA(cudaEventRecord(ctx->binned_ev[slot], ctx->stream));
        CIP_CHECK_CUDA(cudaStreamWaitEvent(gs, ctx->binned_ev[slot], 0));
        const double t0 = now_ms();
        if (timed) CIP_CHECK_CUDA(cudaEventRecord(ctx->chunk_t[ntimed][2], gs));
        CIP_TRY(run_gridder(ctx, slot, gs, g, fp64, g.win_begin, g.win_end - 1, ctx->grid.p, 0, plan->nplanes));
        if (timed) CIP_CHECK_CUDA(cudaEventRecord(ctx->chunk_t[ntimed][3], gs));
        CIP_CHECK_CUDA(cudaEventRecord(ctx->gridded_ev[slot], gs));
        used[slot] = true;
        if (timed) ++ntimed;
        if (rep) rep->nblocks++;
        if (rep && !dev_timing) {
          CIP_CHECK_CUDA(cudaEventSynchronize(ctx->gridded_ev[slot]));
          rep->ms_grid += now_ms() - t0;
        }
        trace_mark("  pass gridded");
      }
    }
    }
    for (int slot = 0; slot < 2; ++slot)
      if (used[slot]) CIP_CHECK_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->gridded_ev[slot], 0));
    if (dev_timing) {
      CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
      float ms = 0;
      CIP_CHECK_CUDA(cudaEventElapsedTime(&ms, ctx->ev[14], ctx->ev[15]));
      rep->ms_zero += ms;  // overlaps the first binning pass
      for (int k = 0; k < ntimed; ++k) {
        CIP_CHECK_CUDA(cudaEventElapsedTime(&ms, ctx->chunk_t[k][0], ctx->chunk_t[k][1]));
        rep->ms_bin += ms;
        CIP_CHECK_CUDA(cudaEventElapsedTime(&ms, ctx->chunk_t[k][2], ctx->chunk_t[k][3]));
        rep->ms_grid += ms;
      }
    }
    tm.start();  // ms_fft includes the w-screen (they alternate per plane)
    CIP_TRY(run_fft_screen(ctx, *plan, ctx->grid.p, 0, plan->nplanes, image_dev, nullptr, nullptr));
    tm.stop(rep ? &rep->ms_fft : &dummy);
    return CIP_OK;
  }

  // Blocks of planes: every block needs all visibilities.
  for (int k = 0; k < nshards; ++k) CIP_TRY(wait_feed(ctx, feeds[k]));
  if (g.W > kFastSupport) {
    // Wide supports have a single window: blocks of resident planes each take a pass over all records (the kernel
    // skips contributions to planes outside the block); one shard is binned once for all blocks.
    if (nshards == 1) {
      tm.start();
      CIP_TRY(run_binning(ctx, 0, in, g, fp64, want_items));
      tm.stop(rep ? &rep->ms_bin : &dummy);
    }
    for (int pb = 0; pb < plan->nplanes; pb += nslots_alloc) {
      const int used = std::min(nslots_alloc, plan->nplanes - pb);
      tm.start();
      CIP_CHECK_CUDA(cudaMemsetAsync(ctx->grid.p, 0, plane_bytes * used, ctx->stream));
      tm.stop(rep ? &rep->ms_zero : &dummy);
      for (int k = 0; k < nshards; ++k) {
        if (nshards > 1) {
          tm.start();
          CIP_TRY(run_binning(ctx, 0, ins[k], g, fp64, want_items));
          tm.stop(rep ? &rep->ms_bin : &dummy);
        }
        tm.start();
        CIP_TRY(run_gridder(ctx, 0, ctx->stream, g, fp64, 0, 0, ctx->grid.p, pb, used));
        tm.stop(rep ? &rep->ms_grid : &dummy);
      }
      tm.start();
      tm.stop(&dummy);
      tm.start();
      CIP_TRY(run_fft_screen(ctx, *plan, ctx->grid.p, pb, used, image_dev, nullptr, nullptr));
      tm.stop(rep ? &rep->ms_fft : &dummy);
      if (rep) rep->nblocks++;
    }
    return CIP_OK;
  }

  // Windows in blocks whose planes fit the stack: one binning pass per block, restricted to its windows.
  const int wblock = std::max(1, std::min(wpp, (nslots_alloc - kWin) / g.win_step + 1));
  for (int wa = 0; wa < g.nwin; wa += wblock) {
    const int wl = std::min(g.nwin, wa + wblock) - 1;
    const int pb = wa * g.win_step;
    const int used = std::min(plan->nplanes, wl * g.win_step + kWin) - pb;
    g.win_begin = wa;
    g.win_end = wl + 1;
    tm.start();
    CIP_CHECK_CUDA(cudaMemsetAsync(ctx->grid.p, 0, plane_bytes * used, ctx->stream));
    tm.stop(rep ? &rep->ms_zero : &dummy);
    for (int k = 0; k < nshards; ++k) {
      tm.start();
      CIP_TRY(run_binning(ctx, 0, ins[k], g, fp64, want_items));
      tm.stop(rep ? &rep->ms_bin : &dummy);
      tm.start();
      CIP_TRY(run_gridder(ctx, 0, ctx->stream, g, fp64, wa, wl, ctx->grid.p, pb, used));
      tm.stop(rep ? &rep->ms_grid : &dummy);
    }
    tm.start();
    CIP_TRY(run_fft_screen(ctx, *plan, ctx->grid.p, pb, used, image_dev, nullptr, nullptr));
    tm.stop(rep ? &rep->ms_fft : &dummy);
    if (rep) rep->nblocks++;
  }
  return CIP_OK;
}

// The binning sizes its buffers from capacities and reports a shortfall through a flag that everything downstream
// honours (nothing touches the image then); the flag is looked at once here, after the attempt has drained.  A
// shortfall grows the buffers to the reported demand and repeats the attempt -- the first call of a new problem
// size at most.
static cip_status grid_accumulate_staged(cip_ctx* ctx, const DevInputs* ins, const VisFeed* feeds, int nshards, cip_plan* plan,
                                         void* image_dev, cip_report* rep) {
  CIP_REQUIRE(plan->support >= kMinSupport && plan->support <= kMaxSupport, "bad support %d", plan->support);
  CIP_REQUIRE(plan->nu >= plan->npix_x && plan->nv >= plan->npix_y && plan->nu % 2 == 0 && plan->nv % 2 == 0,
              "bad grid size %d x %d", plan->nu, plan->nv);
  CIP_REQUIRE(plan->nplanes >= 1 && plan->dw > 0, "bad w-plane layout");
  CIP_REQUIRE(plan->nu >= 2 * kFastSupport + kTile && plan->nv >= 2 * kFastSupport + kTile, "grid too small");
  for (int attempt = 0; attempt < 3; ++attempt) {
    memset(ctx->flags_host, 0, 64);
    CIP_CHECK_CUDA(cudaMemsetAsync(ctx->call_flags.p, 0, 64, ctx->stream));
    cip_report scratch{};
    CIP_TRY(grid_accumulate_attempt(ctx, ins, feeds, nshards, plan, image_dev, rep ? &scratch : nullptr));
    CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    bool again = false;
    CIP_TRY(binning_overflowed(ctx, &again));
    if (!again) {
      if (rep) {
        rep->ms_zero += scratch.ms_zero;
        rep->ms_bin += scratch.ms_bin;
        rep->ms_grid += scratch.ms_grid;
        rep->ms_fft += scratch.ms_fft;
        rep->nblocks += scratch.nblocks;
        if (scratch.ms_h2d > 0) rep->ms_h2d = scratch.ms_h2d;
      }
      return CIP_OK;
    }
    trace_mark("  binning buffers were too small: repeating the pass");
  }
  set_error("binning buffers still too small after two repeats");
  return CIP_ECUDA;
}

extern "C" cip_status cip_grid_accumulate(cip_ctx* ctx, const cip_invert_args* args, cip_plan* plan, void* image_dev,
                                          cip_report* report) {
  CIP_REQUIRE(ctx && plan && image_dev, "NULL argument");
  CIP_TRY(validate_args(args));
  CIP_REQUIRE(plan->npix_x == args->npix_x && plan->npix_y == args->npix_y, "plan and args disagree on the image size");
  CIP_ENTER_DEVICE(ctx);
  if (report) memset(report, 0, sizeof(*report));
  const int64_t l0 = ctx->launches;
  DevInputs in;
  VisFeed feed;
  CopyStreamGuard guard(ctx);
  CIP_TRY(stage_inputs(ctx, args, &in, &feed, true, report ? &report->ms_h2d : nullptr));
  CIP_TRY(grid_accumulate_staged(ctx, &in, &feed, 1, plan, image_dev, report));
  CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
  if (report) report->launches += ctx->launches - l0;
  return CIP_OK;
}

extern "C" cip_status cip_grid_accumulate_many(cip_ctx* ctx, const cip_invert_args* args, int nshards, cip_plan* plan,
                                               void* image_dev, cip_report* report) {
  CIP_REQUIRE(ctx && plan && image_dev && args && nshards >= 1, "NULL argument");
  CIP_ENTER_DEVICE(ctx);
  if (report) memset(report, 0, sizeof(*report));
  const int64_t l0 = ctx->launches;
  std::vector<DevInputs> ins(nshards);
  std::vector<VisFeed> feeds(nshards);
  for (int k = 0; k < nshards; ++k) {
    CIP_TRY(validate_args(&args[k]));
    CIP_REQUIRE(args[k].memspace == 1, "cip_grid_accumulate_many takes device-resident shards (memspace 1)");
    CIP_REQUIRE(plan->npix_x == args[k].npix_x && plan->npix_y == args[k].npix_y, "plan and args disagree on the image size");
    CIP_TRY(stage_inputs(ctx, &args[k], &ins[k], &feeds[k], false, nullptr));
  }
  CIP_TRY(grid_accumulate_staged(ctx, ins.data(), feeds.data(), nshards, plan, image_dev, report));
  CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
  if (report) report->launches += ctx->launches - l0;
  return CIP_OK;
}

extern "C" cip_status cip_bin_inspect(cip_ctx* ctx, const cip_invert_args* args, const cip_plan* plan, int64_t* njobs,
                                      int64_t* ngroups, int* rec_bytes, cip_job* jobs_out, void* recs_out) {
  CIP_REQUIRE(ctx && plan && njobs && ngroups && rec_bytes, "NULL argument");
  static_assert(sizeof(cip_job) == sizeof(Job), "cip_job mirrors the internal job");
  CIP_TRY(validate_args(args));
  CIP_ENTER_DEVICE(ctx);
  DevInputs in;
  VisFeed feed;
  CopyStreamGuard guard(ctx);
  CIP_TRY(stage_inputs(ctx, args, &in, &feed, false, nullptr));
  CIP_TRY(wait_feed(ctx, feed));
  const GridGeom g = make_geom(*plan);
  CIP_REQUIRE(max_windows_per_pass(g) >= g.nwin, "the key table of this plan does not fit one binning pass");
  const int fp64 = plan->accumulate_fp64;
  for (int attempt = 0;; ++attempt) {
    memset(ctx->flags_host, 0, 64);
    CIP_CHECK_CUDA(cudaMemsetAsync(ctx->call_flags.p, 0, 64, ctx->stream));
    CIP_TRY(run_binning(ctx, 0, in, g, fp64, false));
    CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
    bool again = false;
    CIP_TRY(binning_overflowed(ctx, &again));
    if (!again) break;
    CIP_REQUIRE(attempt < 2, "binning buffers still too small after two repeats");
  }
  BinHeader h{};
  CIP_CHECK_CUDA(cudaMemcpy(&h, ctx->work[0].hdr.p, sizeof(h), cudaMemcpyDeviceToHost));
  *njobs = h.njobs;
  *ngroups = h.ngroups;
  *rec_bytes = (!fp64 && g.W <= kFastSupport) ? 16 : 32;
  if (jobs_out && h.njobs) CIP_CHECK_CUDA(cudaMemcpy(jobs_out, ctx->work[0].jobs.p, sizeof(Job) * h.njobs, cudaMemcpyDeviceToHost));
  if (recs_out && h.ngroups)
    CIP_CHECK_CUDA(cudaMemcpy(recs_out, ctx->work[0].recs.p, (size_t)h.ngroups * kGroup * *rec_bytes, cudaMemcpyDeviceToHost));
  return CIP_OK;
}

extern "C" cip_status cip_correct_image(cip_ctx* ctx, const cip_plan* plan, void* image_dev, int row_begin,
                                        int row_end, double scale) {
  CIP_REQUIRE(ctx && plan && image_dev, "NULL argument");
  CIP_ENTER_DEVICE(ctx);
  CIP_TRY(run_correct(ctx, *plan, image_dev, row_begin, row_end, scale));
  CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
  return CIP_OK;
}

extern "C" cip_status cip_invert(cip_ctx* ctx, const cip_invert_args* args, void* dirty_out, cip_scan_result* scan_out,
                                 cip_plan* plan_out, cip_report* report) {
  CIP_REQUIRE(ctx && dirty_out, "NULL argument");
  NvtxRange nvtx("cip_invert");
  CIP_TRY(validate_args(args));
  CIP_ENTER_DEVICE(ctx);
  if (report) memset(report, 0, sizeof(*report));
  const int64_t l0 = ctx->launches;
  cudaStream_t st = ctx->stream;
  trace_mark("begin");
  CIP_CHECK_CUDA(cudaEventRecord(ctx->ev[2], st));
  const int fp64 = args->accumulate_fp64 != 0;
  const size_t img_bytes = (size_t)args->npix_x * args->npix_y * (fp64 ? 8 : 4);
  void* image = dirty_out;
  if (args->memspace == 0) {
    CIP_TRY(ctx->image.ensure(img_bytes));
    image = ctx->image.p;
  }
  CIP_CHECK_CUDA(cudaMemsetAsync(image, 0, img_bytes, st));

  DevInputs in;
  VisFeed feed;
  CopyStreamGuard guard(ctx);
  CIP_TRY(stage_inputs(ctx, args, &in, &feed, false, report ? &report->ms_h2d : nullptr));
  StageTimer tm(ctx, report != nullptr, args->memspace == 0);
  double dummy = 0;
  trace_mark("staged");
  tm.start();
  cip_scan_result scan{};
  CIP_TRY(run_scan(ctx, in, args->do_wstacking, &scan));
  tm.stop(report ? &report->ms_scan : &dummy);
  if (scan_out) *scan_out = scan;
  trace_mark("scanned");
  cip_plan plan{};
  CIP_TRY(cip_plan_make(args->npix_x, args->npix_y, args->pixsize_x, args->pixsize_y, args->epsilon,
                        args->do_wstacking, fp64, scan.wmin, scan.wmax, scan.nactive,
                        getenv("CIP_SIGMA") ? atof(getenv("CIP_SIGMA")) : 0.0,
                        getenv("CIP_SUPPORT") ? atoi(getenv("CIP_SUPPORT")) : 0, &plan));
  trace_mark("planned");
  if (scan.nactive > 0) {
    CIP_TRY(grid_accumulate_staged(ctx, &in, &feed, 1, &plan, image, report));
    trace_mark("accumulated");
    tm.start();
    CIP_TRY(run_correct(ctx, plan, image, 0, args->npix_x, 1.0));
    tm.stop(report ? &report->ms_correct : &dummy);
  }
  if (plan_out) *plan_out = plan;
  if (args->memspace == 0) {
    tm.start();
    CIP_CHECK_CUDA(cudaMemcpyAsync(dirty_out, image, img_bytes, cudaMemcpyDeviceToHost, st));
    tm.stop(report ? &report->ms_d2h : &dummy);
  }
  CIP_CHECK_CUDA(cudaEventRecord(ctx->ev[3], st));
  CIP_CHECK_CUDA(cudaStreamSynchronize(st));
  trace_mark("done");
  if (report) {
    report->ms_total = elapsed_ms(ctx, 2, 3);
    report->launches += ctx->launches - l0;
    report->nactive = scan.nactive;
  }
  return CIP_OK;
}
