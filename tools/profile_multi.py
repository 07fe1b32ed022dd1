#!/usr/bin/env python
"""Development aid: per-phase timing of the multi-rank invert under torchrun (not a test, not the bench).

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 tools/profile_multi.py [--times 64]
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ska-sdp-continuum-imaging-pipeline_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--times", type=int, default=256)
    ap.add_argument("--iters", type=int, default=4)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from ska_sdp_cip import synthetic as S
    from ska_sdp_cip import invert as I

    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = S.ska_mid(n_times=args.times)
    full = S.ska_mid()
    px = S.pixel_size_lm(full, S.make_uvw(full, (0, full.n_times)))
    long_cfg = S.ska_mid(n_times=cfg.n_times * world)
    uvw = S.make_uvw(long_cfg, (rank * cfg.n_times, (rank + 1) * cfg.n_times))
    freq = cfg.frequencies()
    d_vis, d_wgt = S.make_vis_torch(cfg, uvw, freq, px, dev, seed=1234 + rank)
    d_uvw, d_freq = torch.from_numpy(uvw).to(dev), torch.from_numpy(freq).to(dev)
    N = cfg.num_pixels

    def sync():
        torch.cuda.synchronize()

    for it in range(args.iters):
        if world > 1:
            dist.barrier()
        sync()
        t0 = time.perf_counter()
        shard = I.StokesIGridderInput(d_freq, None, d_uvw, d_vis, d_wgt)
        reps = []
        img, plan, sumw = I.accumulate_shards([shard], N, px, reports=reps)
        sync()
        t1 = time.perf_counter()
        mine, r0, r1 = I.reduce_image_rows(img)
        sync()
        t2 = time.perf_counter()
        out = I.reduce_and_correct(img, plan, sumw, normalise=False)
        sync()
        t3 = time.perf_counter()
        r = reps[0] if reps else {}
        print(f"[rank {rank} it {it}] accumulate {1e3 * (t1 - t0):.2f} ms (lib stages: bin {r.get('ms_bin', 0):.2f} "
              f"grid {r.get('ms_grid', 0):.2f} fft {r.get('ms_fft', 0):.2f} zero {r.get('ms_zero', 0):.2f} "
              f"wscr {r.get('ms_wscreen', 0):.2f}) | reduce_rows {1e3 * (t2 - t1):.2f} ms | "
              f"reduce+correct+gather {1e3 * (t3 - t2):.2f} ms", flush=True)
    if world > 1:
        # raw collective timing on a 64 MB image
        x = torch.zeros((N, N), dtype=torch.float32, device=dev)
        y = torch.empty((N // world, N), dtype=torch.float32, device=dev)
        for name, fn in (("reduce_scatter", lambda: dist.reduce_scatter_tensor(y, x)),
                         ("all_gather", lambda: dist.all_gather_into_tensor(x, y)),
                         ("all_reduce", lambda: dist.all_reduce(x))):
            for _ in range(3):
                fn()
            sync()
            t = time.perf_counter()
            for _ in range(10):
                fn()
            sync()
            print(f"[rank {rank}] {name} 64 MB: {1e2 * (time.perf_counter() - t):.3f} ms", flush=True)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
