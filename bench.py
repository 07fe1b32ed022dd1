#!/usr/bin/env python
"""Headline benchmark: invert throughput (Mvis/s) of the B200 hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--times T]

A "step" is one complete invert (scan, binning, gridding, FFTs, w-screen, correction) of one batch
of synthetic SKA-Mid AA* visibilities.  N = 1: BASELINE.json configs[1] (144 dishes = 10,296 baselines x
256 times x 64 channels = 168.7 M visibilities, 4096^2 image, epsilon 1e-4, w-stacking).  N > 1 (launched
under torchrun, one rank per GPU): BASELINE.json configs[3], the stated multi-GPU target -- 10,296
baselines x 512 times x 256 channels = 1.35e9 visibilities, 16384^2 image, sharded by frequency chunk with
the reference's partition(1, freq_chunks=N); every rank grids its channels into a partial image and one
NCCL reduce-scatter / all-gather around the correction combines them: STRONG scaling (the total work is
fixed).  `--workload cfg1|cfg3` overrides the choice (cfg1 at N > 1 is the weak-scaling run of round 1).

Printed JSON (one line, rank 0): `value` = device-resident inputs, timed with CUDA events on the
library's stream; `e2e` = the same step through the reference-facing call
(`ska_sdp_cip.wgridder.ms2dirty` with HOST arrays: pinned host -> device copies and the image read-back
inside the timed region); `roofline` = the gridding kernel against measured HBM bandwidth;
`parity` = relative RMS of the timed path's image against the float64 DFT on the 8 source pixels + 1024
random pixels (each rank transforms its own shard, the sums are all-reduced);
`cpu_baseline` = the oracle's C/OpenMP w-stacking port on a bounded sample, on this box's host cores.
`--impl reference` times that CPU port alone (the reference's own gridder, ducc0, is a third-party
wheel that is not installable here -- see DESIGN.md).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "ska-sdp-continuum-imaging-pipeline_b200"))

METRIC = "invert_throughput"
CPU_SAMPLE_TIMES = 32  # time steps of the workload given to the CPU port (about 10 s on 16 threads)
UNIT = "Mvis/s"


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback"


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one gridding launch, from the committed `ncu --set full`
    capture of this workload (profiles/grid_kernel_traffic.json, written by tools/ncu_summary.py); None if absent."""
    p = ROOT / "profiles" / "grid_kernel_traffic.json"
    try:
        return int(json.loads(p.read_text())["dram_bytes"])
    except Exception:  # noqa: BLE001
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        self.window = None

    def mark(self, t0: float, t1: float):
        """Wall-clock window (time.time()) of the timed region: only samples inside it are reported."""
        self.window = (t0, t1)

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "25"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [l.split(",") for l in Path(self.f.name).read_text().strip().splitlines() if l.count(",") >= 7]
        os.unlink(self.f.name)
        if self.window and rows:
            import datetime
            inside = []
            for r in rows:
                try:
                    ts = datetime.datetime.strptime(r[0].strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
                except ValueError:
                    continue
                if self.window[0] - 0.02 <= ts <= self.window[1] + 0.02:
                    inside.append(r)
            out["samples_total"] = len(rows)
            rows = inside or rows[-3:]  # a region shorter than the sampling period: the samples right around it
        if not rows:
            return out
        sm = [float(r[1]) for r in rows if r[1].strip().replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if r[2].strip().replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows for n, v in zip(names, r[4:8]) if v.strip().lower().startswith("active")})
        out.update(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                   reasons=reasons, samples=len(rows))
        return out


def workload(args, world: int = 1, rank: int = 0):
    """(cfg of the whole job, pixel size, name, uvw of this rank, frequencies of this rank, description)."""
    from ska_sdp_cip import synthetic as S
    from ska_sdp_cip.measurement_set import balanced_chunk_bounds
    name = args.workload or ("cfg1" if world == 1 else "cfg3")
    if name == "cfg3":
        cfg = S.ska_mid(n_times=512, n_chan=256, num_pixels=16384, dfreq_hz=0.5e6)
        uvw = S.make_uvw(cfg)
        px = S.pixel_size_lm(cfg, uvw)
        c0, c1 = list(balanced_chunk_bounds(0, cfg.n_chan, world))[rank]  # the reference's partition(1, world)
        desc = (f"{cfg.name}: {cfg.n_baselines} baselines x {cfg.n_times} times x {cfg.n_chan} channels "
                f"({cfg.n_vis} visibilities in total, {cfg.n_chan // world} channels per GPU), {cfg.num_pixels}^2 image, "
                f"epsilon {cfg.epsilon}, w-stacking (BASELINE.json configs[3])")
        return cfg, px, name, uvw, cfg.frequencies()[c0:c1], desc
    cfg = S.ska_mid(n_times=args.times)
    full = S.ska_mid()  # the pixel size is that of the full 256-time configuration
    px = S.pixel_size_lm(full, S.make_uvw(full, (0, full.n_times)))
    # rank r observes times [r*T, (r+1)*T) of a world*T-step observation (weak scaling)
    long_cfg = S.ska_mid(n_times=cfg.n_times * world)
    uvw = S.make_uvw(long_cfg, (rank * cfg.n_times, (rank + 1) * cfg.n_times))
    desc = (f"{cfg.name}: {cfg.n_baselines} baselines x {cfg.n_times} times x {cfg.n_chan} channels per GPU "
            f"({cfg.n_vis} visibilities), {cfg.num_pixels}^2 image, epsilon {cfg.epsilon}, w-stacking "
            f"(BASELINE.json configs[1])")
    return cfg, px, name, uvw, cfg.frequencies(), desc


_CPU_DATA = {}  # the last sample synthesised for the CPU port (a repeated step times the same arrays)
_CPU_FULL = []  # [(uvw, freq, vis, wgt)] of the whole workload when the caller has them on the host already


def cpu_sample_data(cfg, px, sample_times: int, chan=None):
    """(uvw, freq, vis, wgt) of the first `sample_times` time steps (channels `chan`) of the workload as numpy arrays,
    synthesised in blocks of CPU_SAMPLE_TIMES time steps (bounded temporaries) and kept for the next call."""
    import dataclasses
    from ska_sdp_cip import synthetic as S
    if _CPU_FULL and chan is None:  # time-major rows: the first sample_times steps are the first rows
        uvw, freq, vis, wgt = _CPU_FULL[0]
        n = len(S.make_uvw(cfg, (0, sample_times)))
        return uvw[:n], freq, vis[:n], wgt[:n]
    key = (sample_times, None if chan is None else (chan.start, chan.stop))
    if key not in _CPU_DATA:
        _CPU_DATA.clear()
        freq = cfg.frequencies() if chan is None else cfg.frequencies()[chan]
        blocks = []
        for t0 in range(0, sample_times, CPU_SAMPLE_TIMES):
            t1 = min(sample_times, t0 + CPU_SAMPLE_TIMES)
            uvw = S.make_uvw(cfg, (t0, t1))
            vis, wgt = S.make_vis_numpy(dataclasses.replace(cfg, n_times=t1 - t0), uvw, freq, px, seed=1234 + t0)
            blocks.append((uvw, vis, wgt))
        _CPU_DATA[key] = (np.concatenate([b[0] for b in blocks]), freq, np.concatenate([b[1] for b in blocks]),
                          np.concatenate([b[2] for b in blocks])) if len(blocks) > 1 else (blocks[0][0], freq, blocks[0][1], blocks[0][2])
    return _CPU_DATA[key]


def cpu_port_rate(cfg, px, sample_times: int, chan=None, repeats: int = 1):
    """The CPU arm: the tuned C/OpenMP w-stacking port (oracle/csrc/wgrid_fast.c through oracle.wstack.ms2dirty_fast:
    float32 grids, parallel binning, kernel taps once per record, pocketfft on all threads) on the first
    `sample_times` time steps (channels `chan`) of the workload, full image size.
    Returns (Mvis/s, seconds, visibilities, threads)."""
    from oracle import wstack
    uvw, freq, vis, wgt = cpu_sample_data(cfg, px, sample_times, chan)
    wstack.lib()
    best = None
    for _ in range(repeats):
        t = time.perf_counter()
        wstack.ms2dirty_fast(uvw, freq, vis, wgt, cfg.num_pixels, cfg.num_pixels, px, px, cfg.epsilon)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    nvis = uvw.shape[0] * freq.shape[0]
    return nvis / best / 1e6, best, nvis, wstack.lib().oracle_num_threads()


REFERENCE_BUDGET_S = 240.0  # the whole --impl reference run stays within a few minutes
CPU_FULL_STEP_S = 40.0      # the CPU port takes the WHOLE workload when a step of it is predicted to stay below this


def cpu_full_sample(cfg, wname, sample_times, step_s):
    """Whether the CPU port can be given every time step of the workload instead of the bounded sample: its time grows
    at most linearly with the sample (the transforms and the correction do not grow at all), so step_s * n_times /
    sample_times bounds a full step from above.  configs[1] only (the configs[3] sample is a channel slice)."""
    return wname == "cfg1" and sample_times < cfg.n_times and step_s * cfg.n_times / sample_times <= CPU_FULL_STEP_S



def cpu_sample(cfg, wname, world):
    """Bounded sample of the workload for the CPU port: (time steps, channel slice, description)."""
    from ska_sdp_cip.measurement_set import balanced_chunk_bounds
    if wname == "cfg3":
        c0, c1 = list(balanced_chunk_bounds(0, cfg.n_chan, max(world, 8)))[0]
        return 2, slice(c0, c1), f"first 2 of {cfg.n_times} time steps x channels {c0}..{c1 - 1} of {cfg.n_chan}"
    return CPU_SAMPLE_TIMES, None, f"first {CPU_SAMPLE_TIMES} of {cfg.n_times} time steps, all channels"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg, px, wname, _uvw, _freq, desc = workload(args, args.gpus, 0)
    sample_times, chan, sample_desc = cpu_sample(cfg, wname, args.gpus)
    t_begin = time.perf_counter()
    # warm-up: W steps of the same sample unless one of them already shows that the budget does not allow it
    warm_done = 0
    step_s = None
    for _ in range(args.warmup):
        _, step_s, _, _ = cpu_port_rate(cfg, px, sample_times, chan)
        warm_done += 1
        if warm_done == 1 and cpu_full_sample(cfg, wname, sample_times, step_s):
            # fast enough host: time the whole workload (fixed costs amortised as in the GPU arm's step)
            step_s *= cfg.n_times / sample_times
            sample_times, sample_desc = cfg.n_times, f"all {cfg.n_times} time steps, all channels"
        if step_s * (args.warmup - warm_done + args.steps) > REFERENCE_BUDGET_S:
            break
    left = REFERENCE_BUDGET_S - (time.perf_counter() - t_begin)
    steps = args.steps if step_s is None else max(1, min(args.steps, int(left / max(step_s, 1e-3))))
    rates, t_total, nvis, cores = [], 0.0, 0, 1
    for _ in range(steps):
        r, dt, nvis, cores = cpu_port_rate(cfg, px, sample_times, chan)
        rates.append(r)
        t_total += dt
    value = float(np.mean(rates))
    sample = f"{sample_desc} ({nvis} visibilities) of the same workload, full {cfg.num_pixels}^2 image, per step"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warm_done, "ms_per_step": 1e3 * t_total / max(steps, 1),
            "higher_is_better": True, "scaling": "weak" if wname == "cfg1" else "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc,
                       "note": "ducc0 (the reference's gridder) is not installable here; this is a tuned C/OpenMP "
                               "port of the same algorithm (oracle/csrc/wgrid_fast.c)"
                               + ("" if (steps, warm_done) == (args.steps, args.warmup) else
                                  f"; {steps} timed / {warm_done} warm-up steps instead of {args.steps} / {args.warmup} "
                                  f"to stay within {REFERENCE_BUDGET_S:.0f} s")},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_b200(args):
    import torch
    import torch.distributed as dist
    from ska_sdp_cip import _cabi, synthetic as S, wgridder
    from ska_sdp_cip.invert import (StokesIGridderInput, accumulate_shards, accumulate_shards_by_w, image_to_host,
                                    reduce_and_correct)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    # nvidia-smi needs a second or two before its loop delivers samples: start it before the inputs are synthesised, so
    # that it is sampling by the time the (short) timed region begins
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    cfg, px, wname, uvw, freq, desc = workload(args, world, rank)
    N = cfg.num_pixels
    d_vis, d_wgt = S.make_vis_torch(cfg, uvw, freq, px, dev, seed=1234 + rank)
    d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
    nvis = uvw.shape[0] * freq.shape[0]        # this rank's visibilities
    total_vis = nvis
    if world > 1:
        t = torch.tensor([nvis], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        total_vis = int(t.item())
    ctx = _cabi.Context.get(local)
    stream = torch.cuda.ExternalStream(_cabi.lib().cip_ctx_stream(ctx.handle), device=dev)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    reports = []

    def step_device():
        """One invert with inputs resident in HBM; multi-rank: + NCCL reduction of the partial images."""
        if world == 1:
            img, info = wgridder.ms2dirty(d_uvw, d_freq, d_vis, d_wgt, N, N, px, px, epsilon=cfg.epsilon,
                                          do_wstacking=True, return_info=True)
            reports.append(info["report"])
            return img, info["plan"]
        shard = StokesIGridderInput(d_freq, None, d_uvw, d_vis, d_wgt)
        reps = []
        acc = accumulate_shards_by_w if args.sharding == "w" else accumulate_shards
        t_a = time.perf_counter()
        img, plan, sumw = acc([shard], N, px, epsilon=cfg.epsilon, reports=reps)
        torch.cuda.synchronize()
        t_b = time.perf_counter()  # this rank's partial image is complete: what follows includes the wait for the slowest rank
        img = reduce_and_correct(img, plan, sumw, normalise=False)
        torch.cuda.synchronize()
        if reps:
            reps[-1]["ms_accumulate_host"] = 1e3 * (t_b - t_a)
            reps[-1]["ms_reduce_host"] = 1e3 * (time.perf_counter() - t_b)
        reports.extend(reps)
        return img, plan.as_dict()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(steps):
            flush_buf.zero_()          # flush L2 between iterations (inputs are also >> L2)
            torch.cuda.synchronize()
            t_host = time.perf_counter()
            fn()
            if args.verbose:
                torch.cuda.synchronize()
                print(f"[rank {rank}] step {i}: {1e3 * (time.perf_counter() - t_host):.2f} ms (host clock)",
                      file=sys.stderr, flush=True)
        torch.cuda.current_stream(dev).synchronize()  # collectives run on torch's stream, the library on its own
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(args.warmup):
        step_device()
    reports.clear()
    l0 = ctx_launches(ctx)
    w0 = time.time()
    ms_dev = timed(step_device, args.steps)
    sampler.mark(w0, time.time())
    launches = ctx_launches(ctx) - l0
    clocks = sampler.stop() if rank == 0 else {}
    timed_reports = list(reports)
    img, plan = step_device()

    # ---- parity of the timed path: sampled pixels against the float64 DFT ------------------------------
    parity = None
    if not args.no_parity:
        from oracle import dft, dft_torch
        _, _, _, pi, pj = S.source_list(cfg, px)
        pix = dft_torch.sample_pixels(N, N, 1024, seed=0, always=np.asarray(pi) * N + np.asarray(pj))
        got = img.reshape(-1)[torch.from_numpy(pix).to(dev)].to(torch.float64)
        del img
        torch.cuda.empty_cache()
        ref = torch.from_numpy(dft_torch.dft_pixels(d_uvw, d_freq, d_vis, d_wgt, N, N, px, px, pix, device=dev)).to(dev)
        if world > 1:
            dist.all_reduce(ref)  # the DFT is linear in the samples: the ranks' partial sums add up
        parity = {"rel_rms_dft": dft.rel_rms(got.cpu().numpy(), ref.cpu().numpy()), "pixels": int(len(pix)),
                  "tolerance": 1e-4, "checker": "oracle/dft_torch.py (float64 DFT of all samples of the timed input)"}
    else:
        del img

    # ---- e2e: host buffers through the public call ------------------------------------------
    h_vis = torch.empty(d_vis.shape, dtype=d_vis.dtype, pin_memory=True)
    h_wgt = torch.empty(d_wgt.shape, dtype=d_wgt.dtype, pin_memory=True)
    h_vis.copy_(d_vis)
    h_wgt.copy_(d_wgt)
    n_vis_np, n_wgt_np = h_vis.numpy(), h_wgt.numpy()
    h_uvw, h_freq = torch.from_numpy(uvw).pin_memory(), torch.from_numpy(freq).pin_memory()
    uvw_p, freq_p = h_uvw.numpy(), h_freq.numpy()  # every input of the e2e step sits in pinned host memory
    torch.cuda.synchronize()

    host_reports = []

    def step_host(with_report=False):
        if world == 1:
            # the timed steps do not ask for the stage report: timing the stages of a streamed upload makes the
            # host wait for every chunk (see StageTimer in csrc/invert.cu), which the plain call does not
            if not with_report:
                return wgridder.ms2dirty(uvw_p, freq_p, n_vis_np, n_wgt_np, N, N, px, px, epsilon=cfg.epsilon,
                                         do_wstacking=True)
            img, info = wgridder.ms2dirty(uvw_p, freq_p, n_vis_np, n_wgt_np, N, N, px, px, epsilon=cfg.epsilon,
                                          do_wstacking=True, return_info=True)
            host_reports.append(info["report"])
            return img
        shard = StokesIGridderInput(freq_p, None, uvw_p, n_vis_np, n_wgt_np)
        acc = accumulate_shards_by_w if args.sharding == "w" else accumulate_shards
        img, plan_, sumw = acc([shard], N, px, epsilon=cfg.epsilon)
        return image_to_host(reduce_and_correct(img, plan_, sumw, normalise=False))

    e2e_steps = max(1, min(args.steps, 3))
    step_host()
    host_reports.clear()
    ms_e2e = timed(step_host, e2e_steps)
    if world == 1:
        step_host(with_report=True)  # untimed: stage breakdown of one e2e step, for information
    h2d = int(n_vis_np.nbytes + n_wgt_np.nbytes + uvw.nbytes + freq.nbytes)
    # floor of the e2e step: the same pinned buffers copied host -> device with nothing else going on (all ranks at
    # once: they share the host's memory and PCIe root complexes)
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(e2e_steps):
        d_vis.copy_(h_vis, non_blocking=True)
        d_wgt.copy_(h_wgt, non_blocking=True)
        d_uvw.copy_(h_uvw, non_blocking=True)
    f1.record()
    torch.cuda.synchronize()
    h2d_floor_ms = f0.elapsed_time(f1) / e2e_steps
    if world > 1:
        t = torch.tensor([h2d_floor_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        h2d_floor_ms = float(t.item())
    if world == 1 and wname == "cfg1" and not args.no_cpu:
        _CPU_FULL.append((uvw, freq, n_vis_np, n_wgt_np))  # the CPU leg below times the same host arrays
    del h_vis, h_wgt, n_vis_np, n_wgt_np
    d2h = int(N * N * 4)

    # every rank's share of a sharded step: accumulate (its own work) against reduce (collectives + waiting for the
    # slowest rank), and its planes / stage times -- what shows an unbalanced sharding
    per_rank = None
    if world > 1:
        mine = {"rank": rank,
                "ms_accumulate": float(np.sum([r.get("ms_accumulate_host", 0.0) for r in timed_reports]) / args.steps),
                "ms_reduce": float(np.sum([r.get("ms_reduce_host", 0.0) for r in timed_reports]) / args.steps),
                "nplanes": [r["nplanes"] for r in timed_reports if "nplanes" in r][:1],
                **{k: float(np.sum([r[k] for r in timed_reports]) / args.steps) for k in ("ms_bin", "ms_grid", "ms_fft", "ms_zero")},
                "phases_ms": {k: float(np.sum([r["phases_ms"].get(k, 0.0) for r in timed_reports if "phases_ms" in r]) / args.steps)
                              for k in sorted({k for r in timed_reports for k in r.get("phases_ms", {})})}}
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        per_rank = gathered
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (gridding) ------------------------------------------
    peak, peak_kind = measured_peaks()
    n_grid_launch = sum(r["nblocks"] for r in timed_reports)
    ms_grid = sum(r["ms_grid"] for r in timed_reports) / max(n_grid_launch, 1)
    # algorithmic bytes of one gridding launch: every input read once (12 B + 24 B/row + 8 B/chan) and the
    # plane stack written once (DESIGN.md "kernel 2")
    alg_bytes = nvis * 12 + uvw.shape[0] * 24 + freq.shape[0] * 8 + plan["nplanes"] * plan["nu"] * plan["nv"] * 8
    achieved = alg_bytes / (ms_grid * 1e-3) / 1e9
    stage = {k: float(np.sum([r[k] for r in timed_reports]) / args.steps) for k in
             ("ms_scan", "ms_bin", "ms_zero", "ms_grid", "ms_fft", "ms_wscreen", "ms_correct", "ms_total")}
    phase_keys = sorted({k for r in timed_reports for k in r.get("phases_ms", {})})
    if phase_keys:
        stage["w_sharding_phases_ms"] = {k: float(np.sum([r["phases_ms"].get(k, 0.0) for r in timed_reports
                                                           if "phases_ms" in r]) / args.steps) for k in phase_keys}
        stage["nplanes_of_rank0"] = [r["nplanes"] for r in timed_reports if "nplanes" in r][:4]
    if per_rank is not None:
        stage["per_rank"] = per_rank
    # ---- per-stage rooflines and the whole-invert fraction of SURVEY.md 8(d) --------------------------------
    nrow_, nchan_ = uvw.shape[0], freq.shape[0]
    nu_, nv_, npl_, sg_ = plan["nu"], plan["nv"], plan["nplanes"], plan["sigma"]
    stage_bytes = {
        "ms_scan": nvis * 4 + nrow_ * 24 + nchan_ * 8,                        # weights once, uvw, frequencies
        "ms_bin": nvis * (12 + 16) + 2 * nrow_ * 24,                          # samples read once, 16-byte records written
        "ms_zero": npl_ * nu_ * nv_ * 8,
        "ms_grid": int(alg_bytes),
        "ms_fft": npl_ * (2 * nu_ * nv_ * 8 + 3 * N * nu_ * 8 + 2 * N * N * 4),  # v pass in place, pruned u passes, w-screen
        "ms_correct": 2 * N * N * 4,
    }
    roofline_stages = {k: {"ms": stage[k], "algorithmic_bytes": int(b),
                           "frac": (b / (stage[k] * 1e-3) / 1e9 / peak) if stage[k] > 0 else None}
                       for k, b in stage_bytes.items()}
    b_in = nvis * 12 + nrow_ * 24 + nchan_ * 8
    b_plane = (1 + 4 + 1 / sg_ ** 2) * nu_ * nv_ * 8 + 2 * N * N * 4
    b_alg = b_in + npl_ * b_plane + 2 * N * N * 4
    rate_dev = nvis * args.steps / ms_dev / 1e3 if world == 1 else None       # Mvis/s of one GPU's share
    invert_roofline_mvis = nvis / (b_alg / (peak * 1e9)) / 1e6
    sm_hz = 1e6 * float((clocks or {}).get("sm_mhz") or 1965.0)
    # operand form of the tensor path (grid.cu: two-term bfloat16 where the plan's kernel error is >= 5e-6, else 3xTF32)
    import ctypes as _C
    from ska_sdp_cip import _cabi
    _beta, _err = _C.c_double(), _C.c_double()
    _cabi.lib().cip_kernel_table(float(plan["sigma"]), int(plan["support"]), _C.byref(_beta), _C.byref(_err))
    mma_mode = int(os.environ.get("CIP_GRID_MMA", "2" if _err.value >= 5e-6 else "1"))
    n_mma = {1: 24.0, 2: 16.0}.get(mma_mode)      # mma.sync per group of 8 visibilities (8 tiles x 3 or 2 products)
    k_mma = {1: 8, 2: 16}.get(mma_mode)
    grid_compute = {
        "operands": {0: "FFMA2 rank-1 updates", 1: "3xTF32 (m16n8k8)", 2: "two-term bfloat16 (m16n8k16)"}.get(mma_mode),
        # 8 x 8 x 8 window x (re, im) per visibility: tensor-core flops actually issued for it
        "tensor_tflops": (nvis / 8 * n_mma * 2 * 16 * 8 * k_mma / (stage["ms_grid"] * 1e-3) / 1e12) if n_mma else None,
        # 8 cycles per mma.sync and sub-partition (tools/experiments/mma_rate.cu, both shapes)
        "tensor_pipe_frac": (nvis / 8 * n_mma * 8.0 / (4 * 148 * sm_hz * stage["ms_grid"] * 1e-3)) if n_mma else None,
    }
    # ---- CPU baseline on a bounded sample -----------------------------------------------------
    cpu = None
    if not args.no_cpu and world == 1:  # rank 0 at N = 1 only
        st_, ch_, sd_ = cpu_sample(cfg, wname, world)
        r, dt, nv_s, cores = cpu_port_rate(cfg, px, st_, ch_)
        if cpu_full_sample(cfg, wname, st_, dt):  # fast enough host: the whole workload instead of the sample
            st_, sd_ = cfg.n_times, f"all {cfg.n_times} time steps, all channels"
            r, dt, nv_s, cores = cpu_port_rate(cfg, px, st_, ch_)
        cpu = {"value": r, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{sd_} ({nv_s} visibilities), full {N}^2 image, {dt:.1f} s"}
    line = {
        "metric": METRIC, "value": total_vis * args.steps / ms_dev / 1e3, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "weak" if wname == "cfg1" else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc,
                   "sigma": plan["sigma"], "support": plan["support"], "nu": plan["nu"], "nv": plan["nv"],
                   "nplanes": plan["nplanes"], "l2": "256 MB buffer written between iterations; inputs 2 GB >> L2",
                   "sharding": ("single GPU" if world == 1 else
                                ("frequency chunks of the reference's partition(1, N)" if wname == "cfg3" else
                                 "time ranges per rank")
                                + (", rows exchanged by |w| slab (NCCL all-to-all), one slab per rank" if args.sharding == "w"
                                   else "")
                                + ", NCCL reduce-scatter + all-gather of the dirty image")},
        "parity": parity,
        "e2e": {"value": total_vis * e2e_steps / ms_e2e / 1e3, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "steps": e2e_steps, "ms_per_step": ms_e2e / e2e_steps,
                "h2d_floor_ms": h2d_floor_ms, "h2d_floor_GBps_per_gpu": h2d / h2d_floor_ms / 1e6,
                "stage_ms": {k: float(np.mean([r[k] for r in host_reports])) for k in
                             ("ms_h2d", "ms_scan", "ms_bin", "ms_grid", "ms_fft", "ms_d2h", "ms_total")}
                if host_reports else None},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "grid_warp_kernel<%d>" % mma_mode, "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic(), "peak_source": peak_kind,
                     "ms_per_launch": ms_grid, "algorithmic_bytes": int(alg_bytes), "compute": grid_compute,
                     "invert": {"B_alg_bytes": int(b_alg), "roofline_mvis_per_gpu": invert_roofline_mvis,
                                "frac": (rate_dev / invert_roofline_mvis) if rate_dev else
                                        (nvis / (ms_dev / args.steps * 1e-3) / 1e6 / invert_roofline_mvis),
                                "formula": "SURVEY.md 8(d): B_in + nplanes * ((5 + 1/sigma^2) nu nv 8 + 2 nx ny 4) + 2 nx ny 4"}},
        "roofline_stages": roofline_stages,
        "stage_ms": stage,
        "cpu_baseline": cpu,
        "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def ctx_launches(ctx) -> int:
    """Kernel launches counted by the library (cip_ctx.launches is read through a report delta)."""
    from ska_sdp_cip import _cabi
    return int(_cabi.lib().cip_ctx_launches(ctx.handle))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--times", type=int, default=256, help="time steps per GPU (256 = BASELINE.json configs[1])")
    ap.add_argument("--workload", default=None, choices=["cfg1", "cfg3"],
                    help="cfg1 = BASELINE.json configs[1] per GPU (default at N = 1), cfg3 = configs[3] sharded by "
                         "frequency chunk (default at N > 1)")
    ap.add_argument("--sharding", default="auto", choices=["auto", "freq", "w"],
                    help="N > 1: grid the rank's own frequency chunk (freq), or exchange the rows by |w| slab first so "
                         "that every rank transforms only the planes of its slab (w); auto = w from 4 GPUs on, where the "
                         "halved per-rank FFT work outweighs the exchange (measured on configs[3]: N = 2 216 vs 235 ms, "
                         "N = 8 169 vs 152 ms per step)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the sampled-pixel DFT check of the timed path")
    ap.add_argument("--verbose", action="store_true", help="per-step host timings on stderr")
    args = ap.parse_args()
    if args.sharding == "auto":
        args.sharding = "w" if args.gpus >= 4 else "freq"
    # stdout carries exactly one JSON line: libraries that write to file descriptor 1 behind Python's back (the
    # NCCL version banner) go to stderr while the benchmark runs
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(json_fd, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
