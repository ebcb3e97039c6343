"""Investigation only: host time per call of the N > 1 end-to-end loop (bench.py bench_e2e on host frames), per rank.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/e2e_probe.py \
        --gpus 2 --workload c1 --also '' --no-cpu-baseline --no-parity
Wraps the pipeline's host functions with wall-clock accumulators (no device synchronisation) and prints mean microseconds per
call to stderr when the process ends."""
import atexit
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch.distributed as dist                                           # noqa: E402
from occlusionfusion_b200 import sharded, warpfield                        # noqa: E402

T = {}


def wrap(obj, name, label=None):
    fn = getattr(obj, name)
    label = label or name

    def inner(*a, **k):
        t0 = time.perf_counter()
        r = fn(*a, **k)
        e = T.setdefault(label, [0, 0.0])
        e[0] += 1
        e[1] += time.perf_counter() - t0
        return r
    setattr(obj, name, inner)


for n in ("_stage_host", "_exchange", "pipeline_step", "counters_host", "_integrate", "pipeline_begin", "pipeline_end"):
    wrap(sharded.ShardedFusion, n)
wrap(warpfield.WarpField, "update_transformations")
wrap(dist, "all_gather_into_tensor")


def report():
    rank = int(os.environ.get("RANK", "0"))
    if rank > 1:
        return
    lines = [f"[e2e_probe rank {rank}]"] + [f"  {k:28s} n={v[0]:6d}  mean {1e6 * v[1] / max(1, v[0]):9.1f} us" for k, v in T.items()]
    print("\n".join(lines), file=sys.stderr, flush=True)


atexit.register(report)
import bench                                                               # noqa: E402
sys.exit(bench.main())
