"""GPU, two or more devices: the exchange step (NCCL all-reduce of the fleet totals, issued by the library on its own
communicator) against the CPU oracle over the whole fleet.  Skipped on a one-GPU box; `scripts/nccl_check.py` is the
program the ranks run."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus() -> int:
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_gpus() < 2, reason="needs at least two GPUs")
def test_sharded_rollup_totals_match_the_oracle():
    n = min(_n_gpus(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "scripts", "nccl_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "PASS" in r.stdout


@pytest.mark.skipif(_n_gpus() < 2, reason="needs at least two GPUs")
def test_bench_named_configs_pass_the_allreduce_check():
    """bench.py under torchrun at small sizes: the `configs` block (C3 / C4-strong / C5) carries allreduce_check = pass."""
    import json

    n = min(_n_gpus(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29534", os.path.join(ROOT, "bench.py"), "--gpus", str(n), "--steps", "3", "--warmup", "3", "--nodes", "400",
           "--pods", "40000", "--samples", "5760", "--no-e2e", "--c3-nodes", "1000", "--c3-pods", "100000", "--c5-nodes-per-gpu", "20",
           "--c5-samples", "20000"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["n_gpus"] == n and line["configs"]["allreduce_check"] == "pass", line["configs"]
