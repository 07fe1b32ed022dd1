"""The N > 1 host logic on CPU: two processes over the gloo backend.

What shards: visibilities (time/frequency chunks of the reference's partition); every rank builds a
partial dirty image with a plan all ranks agree on; one reduction over image rows combines them
(SURVEY.md 8e).  The GPU pieces are replaced here by the oracle's w-stacking port so that the
collective plumbing -- shard assignment, statistics agreement, row reduction and gather -- is what
is under test."""
import os
import socket
import traceback

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import random_problem

WORLD = 2


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def worker(rank, world, port, fn, out):
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    for p in (str(root), str(root / "ska-sdp-continuum-imaging-pipeline_b200"), str(root / "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    try:
        dist.init_process_group("gloo", rank=rank, world_size=world)
        out[rank] = fn(rank, world)
    except Exception:  # noqa: BLE001
        out[rank] = "ERR " + traceback.format_exc()
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def run(fn):
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(worker, args=(WORLD, free_port(), fn, out), nprocs=WORLD, join=True)
    res = [out[r] for r in range(WORLD)]
    for r in res:
        assert not (isinstance(r, str) and r.startswith("ERR")), r
    return res


def _stats_case(rank, world):
    from ska_sdp_cip.invert import agree_on_shard_statistics
    local = [(5.0, 100.0, 10, 2.5), (1.0, 50.0, 7, 1.5)][rank]
    return agree_on_shard_statistics(*local)


def test_statistics_agreement_over_gloo():
    res = run(_stats_case)
    assert res[0] == res[1] == (1.0, 100.0, 17, 4.0)


def _empty_rank_case(rank, world):
    from ska_sdp_cip.invert import agree_on_shard_statistics
    local = [(np.inf, 0.0, 0, 0.0), (3.0, 9.0, 4, 2.0)][rank]  # rank 0 holds only flagged data
    return agree_on_shard_statistics(*local)


def test_statistics_agreement_with_an_empty_rank():
    res = run(_empty_rank_case)
    assert res[0] == res[1] == (3.0, 9.0, 4, 2.0)


def _sharded_invert_case(rank, world):
    from oracle import plan as oplan, wstack
    from ska_sdp_cip import _cabi
    from ska_sdp_cip.invert import (agree_on_shard_statistics, gather_image_rows, reduce_image_rows,
                                    shard_of_rank)
    from ska_sdp_cip.measurement_set import balanced_chunk_bounds
    uvw, freq, vis, wgt, nx, px = random_problem(nrow=600, nchan=6, nx=64)
    # the reference's partition: 2 row chunks x 3 frequency chunks, row-major / frequency-minor
    chunks = [(r, c) for r in balanced_chunk_bounds(0, 600, 2) for c in balanced_chunk_bounds(0, 6, 3)]
    mine = shard_of_rank(chunks, rank, world)
    stats = [np.inf, 0.0, 0, 0.0]
    for (r0, r1), (c0, c1) in mine:
        a, b, n, s = wstack.scan(uvw[r0:r1], freq[c0:c1], np.ascontiguousarray(vis[r0:r1, c0:c1]),
                                 np.ascontiguousarray(wgt[r0:r1, c0:c1]))
        stats = [min(stats[0], a), max(stats[1], b), stats[2] + n, stats[3] + s]
    wmin, wmax, nact, sumw = agree_on_shard_statistics(*stats)
    p = _cabi.plan_make(nx, nx, px, px, 1e-4, wmin, wmax, nact, fp64=True)  # host-only C-ABI call
    op = oplan.make_plan(nx, nx, px, px, 1e-4, wmin, wmax, nact, fp64=True)
    assert (p.sigma, p.support, p.nplanes) == (op.sigma, op.W, op.nplanes)
    partial = np.zeros((nx, nx))
    for (r0, r1), (c0, c1) in mine:
        partial += wstack.ms2dirty(uvw[r0:r1], freq[c0:c1], np.ascontiguousarray(vis[r0:r1, c0:c1]),
                                   np.ascontiguousarray(wgt[r0:r1, c0:c1]), nx, nx, px, px, 1e-4, fp64=True, plan=op)
    image = torch.from_numpy(partial)
    rows, b, e = reduce_image_rows(image)
    assert (b, e) == (rank * nx // world, (rank + 1) * nx // world) and rows.shape == (nx // world, nx)
    rows = rows * (1.0 / sumw)  # stands in for cip_correct_image(scale = 1 / sum of weights) on this row block
    full = gather_image_rows(image, rows, b, e)
    return full.numpy().copy(), sumw, len(mine), p.as_dict()


def test_sharded_invert_over_gloo_matches_the_unsharded_image():
    from oracle import dft
    res = run(_sharded_invert_case)
    uvw, freq, vis, wgt, nx, px = random_problem(nrow=600, nchan=6, nx=64)
    (img0, sumw0, n0, plan0), (img1, sumw1, n1, plan1) = res
    assert n0 == n1 == 3 and plan0 == plan1
    assert sumw0 == sumw1 == pytest.approx(float(wgt.sum(dtype=np.float64)), rel=1e-6)
    assert np.array_equal(img0, img1)  # every rank ends with the same full image
    ref = dft.dft_dirty(uvw, freq, vis, wgt, nx, nx, px, px) / wgt.sum(dtype=np.float64)
    assert dft.rel_rms(img0, ref) <= 1e-5


def _odd_rows_case(rank, world):
    from ska_sdp_cip.invert import gather_image_rows, reduce_image_rows
    image = torch.full((5, 4), float(rank + 1), dtype=torch.float64)  # 5 rows do not divide by 2
    rows, b, e = reduce_image_rows(image)
    full = gather_image_rows(image, rows, b, e)
    return (b, e), full.numpy().copy()


def test_row_reduction_falls_back_to_all_reduce_for_indivisible_images():
    for (b, e), full in run(_odd_rows_case):
        assert (b, e) == (0, 5) and np.all(full == 3.0)


def test_shard_assignment_is_a_partition():
    from ska_sdp_cip.invert import shard_of_rank
    chunks = list(range(11))
    for world in (1, 2, 3, 8, 16):
        parts = [shard_of_rank(chunks, r, world) for r in range(world)]
        assert sorted(sum(parts, [])) == chunks
        assert max(map(len, parts)) - min(map(len, parts)) <= 1
