"""GPU, two ranks (skips itself on a box with fewer than two GPUs): bench.py under torch.distributed.run exactly as the
driver launches it -- one process per GPU over NCCL, the pipelined transport, and the post-run parity check of EVERY
rank's shard against the CPU oracle through the same driver (bench.py parity_check).  A mismatch on any rank makes
bench.py exit non-zero."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run_bench(n, extra, timeout=900, env=None):
    import torch
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs, this box has {torch.cuda.device_count()}")
    port = 29700 + os.getpid() % 200
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", str(n)] + extra
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout, env={**os.environ, **(env or {})})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout[-2000:]
    return json.loads(lines[0])


def test_two_ranks_pipelined_parity_through_bench():
    rec = _run_bench(2, ["--steps", "4", "--warmup", "3", "--workload", "c1", "--also", "", "--no-cpu-baseline"])
    assert rec["n_gpus"] == 2 and rec["config"]["workload"].startswith("c1")
    assert rec["parity"]["checked"] > 100_000 and rec["parity"]["mismatches"] == 0
    assert rec["gpu_launches"] > 0 and rec["value"] > 0 and rec["e2e"]["value"] > 0


def test_two_ranks_sequential_transport_parity():
    rec = _run_bench(2, ["--steps", "3", "--warmup", "3", "--workload", "c1", "--also", "", "--no-cpu-baseline", "--no-pipeline",
                         "--no-e2e"])
    assert rec["parity"]["mismatches"] == 0 and rec["parity"]["checked"] > 100_000


def test_two_ranks_multicast_push_parity():
    """The packet pushed once through the NVSwitch multicast mapping (ofx_multicast_copy + a symmetric-memory barrier)
    instead of being pulled by every peer: forced on for a small packet (the default uses it from 4 MB)."""
    rec = _run_bench(2, ["--steps", "4", "--warmup", "3", "--workload", "c1", "--also", "", "--no-cpu-baseline", "--no-e2e"],
                     env={"OFX_MULTICAST": "1", "OFX_DEBUG": "1"})
    assert rec["parity"]["mismatches"] == 0 and rec["parity"]["checked"] > 100_000
