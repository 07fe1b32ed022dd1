"""The real NCCL path: a 2-rank torchrun of tests/nccl_parity_main.py.  Needs two GPUs;
on a one-GPU box the test is skipped (the committed log profiles/r2_nccl_parity_n2.json holds a 2-GPU run)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu


def test_two_rank_nccl_reduce_matches_single_gpu_and_dft(built_lib):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = Path(__file__).resolve().parent / "nccl_parity_main.py"
    env = dict(os.environ, OMP_NUM_THREADS="4")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29631", str(script)],
                         capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    line = [l for l in out.stdout.splitlines() if l.startswith('{"nccl_parity"')][-1]
    res = json.loads(line)
    assert res["ok"] and res["nccl_parity"]["even"]["world"] == 2
