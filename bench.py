#!/usr/bin/env python
"""bench.py -- warped-TSDF integration throughput (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2-dense] [--impl reference]

A step = one frame: node-transform upload + fused warp/integrate of the whole volume.
  value      updated voxels / s, inputs resident in HBM, CUDA-event time, max over ranks
  e2e        the same through the public classes (WarpField.update_transformations +
             TSDFVolume.integrate on HOST numpy buffers, + reading the counters back)
  roofline   the fused kernel against the measured HBM copy bandwidth (MEASURED_PEAKS.json)
  cpu_baseline  the CPU oracle (C/OpenMP port of the reference CPU path) on a bounded slab
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "warped_tsdf_voxel_updates_per_sec"
UNIT = "voxel-updates/s"
FALLBACK_HBM_GBS = 6650.0     # /opt/skills/guides/B200_PROFILING.md fallback


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c2-dense")
    ap.add_argument("--impl", default="ofx", choices=["ofx", "reference"])
    ap.add_argument("--also", default="c2,c1", help="extra workloads reported under 'also' (N=1 only); '' = none")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA graph replay")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (investigation runs)")
    ap.add_argument("--no-pipeline", action="store_true", help="N > 1: broadcast -> kernel -> all-gather strictly in sequence")
    ap.add_argument("--cpu-planes", type=int, default=0, help="x-planes in the CPU sample (0 = auto)")
    ap.add_argument("--graph-ops", action="store_true",
                    help="time the graph-maintenance / surface-mask operators (SURVEY section 8 rows f2/f3) instead of the "
                         "integrate path: one JSON line, each operator next to its CPU baseline")
    return ap.parse_args()


def ncu_traffic(workload):
    """DRAM bytes (read + write) of ONE launch of the fused kernel on this workload, from the committed
    `ncu --set full` capture (profiles/ncu_traffic.json: written by hand from the .ncu-rep, per round)."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        return json.load(open(p)).get(workload)
    except Exception:
        return None


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
def cpu_baseline(w, n_steps, n_planes=0, n_frames=4):
    """The CPU oracle (port of the reference CPU path: cached skin -> deform -> integrate CPU
    branch) on a bounded sample: a centred slab of x-planes of the SAME volume, steady state
    (skin cached, as WarpField.skin_tsdf does), all host cores via OpenMP."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    O.build()
    cores = os.cpu_count() or 1
    Dx, Dy, Dz = w.dims
    if n_planes <= 0:
        n_planes = max(1, min(Dx, int(6e6 // (Dy * Dz)) or 1))      # ~6 M voxels: seconds of CPU work
    x0 = (Dx - n_planes) // 2
    ov = O.OracleVolume((n_planes, Dy, Dz), w.origin, w.voxel_size, w.intr, w.trunc, x0=x0)
    ov.set_graph(w.nodes, w.coverage, w.K)
    ov.integrate(w.frame(0), True)
    t0 = time.perf_counter()
    ov.skin_tsdf()
    t_skin = time.perf_counter() - t0
    frames = [w.frame(1 + i) for i in range(n_frames)]
    ests = [w.estimate(1 + i) for i in range(n_frames)]
    Ts = [O.origin_translations(e["node_rotations"], e["node_translations"], w.nodes) for e in ests]
    def one(i):
        t0 = time.perf_counter()
        ov.set_transforms(ests[i]["node_rotations"], Ts[i])
        n = ov.integrate(frames[i], False)
        return time.perf_counter() - t0, n

    # pick the thread count that is fastest on this host (more threads is not always faster:
    # cgroup-limited containers and memory-bound loops), then time n_steps frames with it
    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else cores
    cands = sorted({1, 2, 4, 8, 16, 32, avail // 2, avail} & set(range(1, avail + 1)))
    trial = {}
    one(0)
    for nt in cands:
        O.set_threads(nt)
        one(0)
        trial[nt] = min(one(1)[0], one(2)[0])
    best = min(trial, key=trial.get)
    O.set_threads(best)
    cores = best
    times, upd = [], []
    for s in range(n_steps):
        dt, n = one(s % n_frames)
        times.append(dt); upd.append(n)
    t = float(np.mean(times))
    n_upd = float(np.mean(upd))
    return {"value": n_upd / t, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle (C/OpenMP port of tsdf.py:442-494 + warpfield.py deform, cached skin) on x-planes "
                      f"[{x0},{x0 + n_planes}) of {w.name} {Dx}x{Dy}x{Dz} = {n_planes * Dy * Dz} voxels, "
                      f"{n_steps} frames, {n_upd:.0f} updates/frame, {t * 1e3:.1f} ms/frame "
                      f"(one-off KNN skin of the slab: {t_skin:.1f} s, not in value); OpenMP threads tried "
                      f"{ {k: round(v * 1e3, 1) for k, v in trial.items()} } ms -> {best} used of {avail} available",
            "ms_per_frame_sample": t * 1e3, "voxels_visited_per_s": n_planes * Dy * Dz / t}


# ------------------------------------------------------------------------------------------------
def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (here the oracle port:
    the reference is Python and /root/reference does not exist on the GPU box), all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from occlusionfusion_b200.synthetic import Workload
    w = Workload(args.workload)
    cb = cpu_baseline(w, max(1, args.steps), args.cpu_planes)
    ms = cb["ms_per_frame_sample"]
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (f64 projection/TSDF average)", "data": "synthetic",
            "config": workload_config(w, args.gpus), "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(w, n_gpus, pipelined=False):
    return {"workload": f"{w.name}: {w.dims[0]}x{w.dims[1]}x{w.dims[2]} voxels @ {w.voxel_size * 1e3:.2f} mm, "
                        f"{w.W}x{w.H} synthetic depth+colour, {w.M} nodes, K={w.K}, scene={w.scene}, warped frames, "
                        f"steady state (graph unchanged)",
            "sharding": f"x-plane blocks of 4 dealt round-robin over {n_gpus} GPU(s)" if n_gpus > 1 else "single GPU",
            "l2": ("per-GPU working set per frame (volume words + skin cache, >= 300 MB) exceeds the 126 MB L2; "
                   "no flush between pipelined steps") if pipelined else
                  "volume state > L2 and a 256 MiB buffer is written between timed steps",
            "launch": ("two packets / two streams: broadcast of frame t+1 and counter all-gather of frame t overlap the "
                       "kernels; the K steps are one CUDA graph (falls back to eager launches), timed first-to-last with CUDA "
                       "events, max over ranks") if pipelined else
                      "per-frame launch sequence replayed as one CUDA graph (falls back to eager launches)"}


# ------------------------------------------------------------------------------------------------
def bench_workload(name, args, rank, world, device, with_e2e=True):
    """Returns a dict of measurements for one workload on this rank's shard (max over ranks applied)."""
    import torch
    import torch.distributed as dist
    from occlusionfusion_b200 import TSDFVolume, WarpField, _capi
    from occlusionfusion_b200.sharded import CudaShard, ShardedFusion, default_x_block
    from occlusionfusion_b200.synthetic import Workload

    w = Workload(name)
    xb = default_x_block(w.dims[0], world)
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None, device=device, shard=rank, n_shards=world, x_block=xb)
    wf = WarpField(w.graph(), vol, None)
    vol.warpfield = wf
    sf = ShardedFusion(CudaShard(vol), w.dims, rank, world, xb, device)
    n_frames = 4
    frames = [w.frame(i) for i in range(n_frames + 1)] if rank == 0 else [None] * (n_frames + 1)
    ests = [w.estimate(i) for i in range(1, n_frames + 1)]
    H, W = w.H, w.W

    # source frame + graph
    if rank == 0:
        sf._ensure_packet(H, W)
    sf.set_graph(w.nodes if rank == 0 else None, w.coverage, w.K)
    if rank == 0:
        sf.fill_packet(frames[0], np.tile(np.eye(3, dtype=np.float32), (w.M, 1, 1)), np.zeros((w.M, 3), np.float32))
    sf.step(H, W, identity_warp=True)
    wf._graph_key = (id(wf.graph.nodes), w.M, w.K, float(w.coverage))      # graph already on the device
    n_active = None

    # device-resident frame packets (inputs in HBM before the timed region)
    packets = []
    for i in range(n_frames):
        if rank == 0:
            wf.rotations = ests[i]["node_rotations"]
            T = -np.einsum("nij,nj->ni", ests[i]["node_rotations"], w.nodes) + w.nodes + ests[i]["node_translations"]
            sf.fill_packet(frames[i + 1], ests[i]["node_rotations"], T)
        torch.cuda.synchronize(device)
        packets.append(sf.buf.clone())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    stream = torch.cuda.current_stream(device)

    gath = torch.empty(world * 4, dtype=torch.int64, device=device)

    def core(kev0=None, kev1=None):
        """the per-frame device work: [broadcast] -> frame preparation (+ node transforms) -> brick cull -> fused kernel -> [counter all-gather]"""
        if world > 1:
            dist.broadcast(sf.buf, src=0)
        v = sf.views
        if kev0 is not None:
            kev0.record(torch.cuda.current_stream(device))
        st = sf.be.integrate_warped(v["R"], v["T"], v["depth"], v["colour"], 1.0)
        if kev1 is not None:
            kev1.record(torch.cuda.current_stream(device))
        if world > 1:
            dist.all_gather_into_tensor(gath, st)
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    def prep(i):
        flush.fill_(i & 0xFF)                                  # evict the volume lines from L2
        sf.buf.copy_(packets[i % n_frames])                    # this frame's packet (device copy, untimed)

    mk = lambda: torch.cuda.Event(enable_timing=True)
    for i in range(max(3, args.warmup)):
        prep(i)
        core()
    barrier()
    K = args.steps

    # pass A (eager launches): kernel-only duration of the fused kernel + the counters
    kevs = [(mk(), mk()) for _ in range(K)]
    stats_acc = torch.zeros(4, dtype=torch.int64, device=device)
    for i in range(K):
        prep(i)
        stats_acc += core(*kevs[i])
    barrier()

    pipelined = world > 1 and not args.no_pipeline
    graph = None
    if pipelined:
        # ---- pass B (N > 1): two packets, two streams -- frame t+1's broadcast and frame t's counter
        # all-gather overlap the kernels (sharded.py: pipeline_*).  Timed first-to-last on the compute
        # stream; no L2 flush: a GPU's per-frame working set (volume words + skin cache, > 300 MB at
        # N = 8) exceeds the 126 MB L2.
        sampler = ClockSampler(device.index or 0)
        for i in range(3):                                     # warm the second packet / stream / NCCL channels
            sf.pipeline_begin(H, W, packets[0] if rank == 0 else None)
            sf.pipeline_step(packets[1] if rank == 0 else None)
            sf.pipeline_step(last=True)
            sf.pipeline_end()
        def run_pipeline():
            sf.pipeline_begin(H, W, packets[0] if rank == 0 else None)
            for i in range(K):
                nxt = packets[(i + 1) % n_frames] if rank == 0 else None
                sf.pipeline_step(nxt, last=(i == K - 1))
            sf.pipeline_end()

        # the K-step launch sequence (both streams, NCCL included) as one CUDA graph: at N = 8 a step is
        # ~0.1 ms on the device, less than Python needs to enqueue its ~10 launches
        if not args.no_graph:
            try:
                g_ = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g_):
                    run_pipeline()
                graph = g_
                graph.replay()
                barrier()
            except Exception as e:            # noqa: BLE001
                print(f"[bench] CUDA graph capture unavailable ({type(e).__name__}: {e}); timing eager launches", file=sys.stderr)
                graph = None
                barrier()
        if rank == 0:
            sampler.start()
        barrier()
        e0, e1 = mk(), mk()
        t_wall0 = time.perf_counter()
        e0.record(stream)
        if graph is not None:
            graph.replay()
        else:
            run_pipeline()
        e1.record(stream)
        barrier()
        t_wall = time.perf_counter() - t_wall0
        clocks = sampler.stop() if rank == 0 else None
        step_ms_total = e0.elapsed_time(e1)
        evs = [(None, None)] * K
    else:
        # the per-frame launch sequence is short and launch-bound: replay it as one CUDA graph
        if not args.no_graph:
            try:
                g_ = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g_):
                    core()
                graph = g_
                prep(0); graph.replay()
                barrier()
            except Exception as e:            # noqa: BLE001
                print(f"[bench] CUDA graph capture unavailable ({type(e).__name__}: {e}); timing eager launches", file=sys.stderr)
                graph = None
                barrier()

        # pass B: the timed K steps
        evs = [(mk(), mk()) for _ in range(K)]
        sampler = ClockSampler(device.index or 0)
        if rank == 0:
            sampler.start()
        barrier()
        t_wall0 = time.perf_counter()
        for i in range(K):
            prep(i)
            evs[i][0].record(stream)
            if graph is not None:
                graph.replay()
            else:
                core()
            evs[i][1].record(stream)
        barrier()
        t_wall = time.perf_counter() - t_wall0
        clocks = sampler.stop() if rank == 0 else None
        step_ms_total = sum(a.elapsed_time(b) for a, b in evs)
    evs = [(a_, b_, c_, d_) for (a_, b_), (c_, d_) in zip(evs, kevs)]
    step_ms = step_ms_total
    kern_ms = sum(c.elapsed_time(d) for _, _, c, d in evs)
    tot = torch.tensor([step_ms, kern_ms], dtype=torch.float64, device=device)
    cnt = stats_acc.clone()
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    step_ms, kern_ms = float(tot[0]) / K, float(tot[1]) / K
    upd = float(cnt[_capi.STAT_UPDATED]) / K
    valid = float(cnt[_capi.STAT_SKIN_VALID]) / K
    bricks = float(cnt[_capi.STAT_BRICKS]) / K
    local_upd = float(stats_acc[_capi.STAT_UPDATED]) / K
    res = dict(workload=w, ms_per_step=step_ms, kernel_ms=kern_ms, updated=upd, skin_valid=valid, bricks=bricks,
               value=upd / (step_ms * 1e-3), wall_ms_per_step=t_wall * 1e3 / K, clocks=clocks,
               local_updated=local_upd, launches_per_step=3, graph=graph is not None, pipelined=pipelined)   # node transforms, brick cull, fused kernel
    # roofline of the fused kernel on this rank's shard (rank 0 reports its own shard)
    alg_bytes = 24.0 * local_upd + 8.0 * H * W + 60.0 * w.M
    res["alg_bytes"] = alg_bytes
    res["achieved_gbs"] = alg_bytes / (float(kern_ms_local(evs, K)) * 1e-3) / 1e9
    res["kernel_ms_local"] = kern_ms_local(evs, K)

    # ---- e2e: public API, host buffers, H2D + D2H inside the timed region ------------------------
    if with_e2e and not args.no_e2e:
        res["e2e"] = bench_e2e(w, vol, wf, sf, frames, ests, args, rank, world, device)
    del vol, wf, sf, packets, flush
    torch.cuda.empty_cache()
    return res


def kern_ms_local(evs, K):
    return sum(c.elapsed_time(d) for _, _, c, d in evs) / K


def bench_e2e(w, vol, wf, sf, frames, ests, args, rank, world, device):
    """Per step, through the public classes: rank 0 takes the HOST frame + the estimator dict,
    WarpField.update_transformations (host numpy), H2D of RGB+Z (pinned staging) and of R,T;
    [N>1: one NCCL broadcast]; fused kernel; counters read back to the host."""
    import torch
    import torch.distributed as dist
    n_frames = len(ests)
    K, Wm = args.steps, max(3, args.warmup)
    H, W = w.H, w.W
    # the camera frames wait in page-locked host memory (numpy views of pinned tensors): the H2D copy of
    # RGB + Z inside the timed region is a DMA straight from them
    if rank == 0:
        frames = [torch.from_numpy(f).pin_memory().numpy() if f is not None else None for f in frames]
    fid = [getattr(vol, "frame_id", 0)]
    wf.frame_id = fid[0]                 # the warp field follows the volume's frame counter (warpfield.py:399)

    def step(i):
        est = dict(ests[i % n_frames])
        fid[0] += 1
        est["target_frame_id"] = fid[0]
        if world == 1:
            wf.update_transformations(est)
            vol.integrate({"im": frames[1 + i % n_frames], "id": fid[0]})
            return vol.last_stats()["updated"]
        if rank == 0:
            wf.frame_id = vol.frame_id
            wf.update_transformations(est)
            sf.fill_packet(frames[1 + i % n_frames], wf.rotations, wf.translations)
        else:
            vol.frame_id = fid[0]
        out = sf.step(H, W, identity_warp=False)
        return int(out[:, 0].sum().item())

    for i in range(Wm):
        step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    upd = 0
    for i in range(K):
        upd += step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    dt = time.perf_counter() - t0
    tmax = torch.tensor([dt], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dt = float(tmax[0])
    h2d = 4 * 4 * H * W + 48 * w.M
    return {"value": upd / dt, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 32 * world,
            "ms_per_step": dt * 1e3 / K, "timing": "host wall clock around K steps (includes host numpy + sync per step)"}


# ------------------------------------------------------------------------------------------------
def bench_graph_ops(args, device):
    """Rows f2 / f3 (ofx_graph.cu), each operator timed with CUDA events (median of 5 after a warm-up call, inputs resident)
    next to its CPU baseline on the same input: the reference's own C++ (oracle/_ref/ref_graph.so, kind "reference") for
    erode_mesh / sample_nodes when it loads, else the oracle port; the oracle port (C, OpenMP where the reference's numba
    prange is parallel) for the rest.  truncated_region also carries a roofline (5 B/voxel: 4 read + 1 written)."""
    import ctypes
    import torch
    from occlusionfusion_b200 import _capi, TSDFVolume
    from occlusionfusion_b200.synthetic import Workload
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    O.build()
    ref_graph = None
    try:
        sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
        import ref_graph
    except Exception as e:      # noqa: BLE001
        print(f"bench: oracle/_ref/ref_graph.so not loadable ({e}); CPU baselines use the oracle port", file=sys.stderr)
    lib = _capi.load()
    S = lambda: _capi.stream_ptr(device)      # noqa: E731

    def gpu_ms(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    def cpu_ms(fn, reps=3):
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            r = fn()
            ts.append((time.perf_counter() - t0) * 1e3)
        return float(np.median(ts)), r

    def mesh(n, rng):
        u, v = np.meshgrid(np.linspace(-0.8, 0.8, n), np.linspace(-0.6, 0.6, n), indexing="ij")
        z = 1.0 + 0.2 * np.sin(4 * u) * np.cos(3 * v)
        verts = (np.stack([u, v, z], -1).reshape(-1, 3) + rng.uniform(-1e-3, 1e-3, (n * n, 3))).astype(np.float32)
        ij = np.argwhere(rng.random((n - 1, n - 1)) > 0.05)
        a = ij[:, 0] * n + ij[:, 1]
        faces = np.concatenate([np.stack([a, a + 1, a + n], 1), np.stack([a + 1, a + n + 1, a + n], 1)]).astype(np.int32)
        return verts, faces

    peak, peak_src = hbm_peak()
    cores = O.max_threads()
    out = {"metric": "graph_ops_ms", "unit": "ms", "higher_is_better": False, "n_gpus": 1, "data": "synthetic",
           "gpu": torch.cuda.get_device_name(device), "host_cores": os.cpu_count(), "ops": {}}
    rng = np.random.default_rng(11)
    cov = 0.05
    for n in (300, 700):
        verts, faces = mesh(n, rng)
        V, F = len(verts), len(faces)
        v_d, f_d = torch.from_numpy(verts).to(device), torch.from_numpy(faces).to(device)
        ne = torch.empty(V, dtype=torch.uint8, device=device)
        t_erode = gpu_ms(lambda: _capi.check(lib.ofx_erode_mesh(_capi.ptr(f_d), F, V, 3, 3, _capi.ptr(ne), S())))
        if ref_graph is not None:
            c_erode, ne_ref = cpu_ms(lambda: np.asarray(ref_graph.erode_mesh(verts, faces, 3, 3)).reshape(-1))
        else:
            c_erode, ne_ref = cpu_ms(lambda: O.erode_mesh(V, faces, 3, 3))
        assert np.array_equal(ne.cpu().numpy().astype(bool), ne_ref)
        idx = torch.empty(V, dtype=torch.int32, device=device)
        pos = torch.empty((V, 3), dtype=torch.float32, device=device)
        cnt = torch.zeros(1, dtype=torch.int32, device=device)
        t_sample = gpu_ms(lambda: _capi.check(lib.ofx_sample_nodes(_capi.ptr(v_d), V, _capi.ptr(ne), ctypes.c_float(cov), 1,
                                                                   _capi.ptr(idx), _capi.ptr(pos), _capi.ptr(cnt), S())), reps=3)
        M = int(cnt.item())
        if ref_graph is not None:
            def ref_sample():
                p_, i_ = np.zeros(0, np.float32), np.zeros(0, np.int32)
                k = ref_graph.sample_nodes(verts, ne_ref.reshape(-1, 1), p_, i_, cov, True, False)
                return i_[:k].reshape(-1).copy()
            c_sample, idx_ref = cpu_ms(ref_sample, reps=1)
        else:
            c_sample, (_, idx_ref) = cpu_ms(lambda: O.sample_nodes(verts, ne_ref, cov, True), reps=1)
        assert np.array_equal(idx[:M].cpu().numpy(), idx_ref)
        nodes = pos[: M // 2].contiguous()                       # a graph that covers about half of the mesh
        nodes_h = nodes.cpu().numpy()
        dist = torch.empty(V, dtype=torch.float32, device=device)
        t_near = gpu_ms(lambda: _capi.check(lib.ofx_nearest_node(_capi.ptr(v_d), V, _capi.ptr(nodes), nodes.shape[0],
                                                                 _capi.ptr(dist), None, S())))
        c_near, (d_ref, _) = cpu_ms(lambda: O.nearest_node(verts, nodes_h), reps=1)
        d = dist.cpu().numpy()
        assert np.array_equal(d, d_ref)
        cand = np.where(d > 2 * cov)[0]
        cand = cand[np.argsort(d[cand])[::-1]].astype(np.int32)
        c_d = torch.from_numpy(cand).to(device)
        acc = torch.empty(max(len(cand), 1), dtype=torch.int32, device=device)
        t_greedy = gpu_ms(lambda: _capi.check(lib.ofx_greedy_insert(_capi.ptr(v_d), V, _capi.ptr(c_d), len(cand), _capi.ptr(nodes),
                                                                    nodes.shape[0], cov, _capi.ptr(acc), _capi.ptr(cnt), S())), reps=3)
        c_greedy, acc_ref = cpu_ms(lambda: O.greedy_update(verts, cand, nodes_h, cov), reps=1)
        k = int(cnt.item())
        assert np.array_equal(acc[:k].cpu().numpy(), acc_ref)
        kind = "reference" if ref_graph is not None else "port"
        # host half of the builder: geodesic edges between the sampled nodes (host threads) next to the reference's serial C++
        from occlusionfusion_b200.graph_proc import compute_edges_geodesic_c
        nidx = idx[:M].cpu().numpy().reshape(-1, 1)
        ge, gw, gd = -np.ones((M, 8), np.int32), np.zeros((M, 8), np.float32), np.zeros((M, 8), np.float32)
        t0 = time.perf_counter()
        compute_edges_geodesic_c(verts, np.ones(V, bool), faces, nidx, 8, cov, ge, gw, gd, None, True, True)
        t_geo = (time.perf_counter() - t0) * 1e3
        c_geo = None
        if ref_graph is not None and V <= 100000:           # the reference needs a dense (M, V) float32 scratch matrix
            re_, rw, rd = -np.ones((M, 8), np.int32), np.zeros((M, 8), np.float32), np.zeros((M, 8), np.float32)
            n2v = -np.ones((M, V), np.float32)
            t0 = time.perf_counter()
            ref_graph.compute_edges_geodesic(verts, np.ones((V, 1), bool), faces, nidx, 8, cov, re_, rw, rd, n2v, True, True)
            c_geo = (time.perf_counter() - t0) * 1e3
            assert np.array_equal(ge, re_) and np.array_equal(gw, rw) and np.array_equal(gd, rd)
        out["ops"][f"mesh_{n}x{n}"] = {
            "edges_geodesic_host": {"ms": t_geo, "host_threads": os.cpu_count(), "cpu_ms": c_geo, "cpu_kind": "reference",
                                    "cpu_cores": 1, "where": "host C++ in libofx (the reference's is host C++ too)"},
            "V": V, "F": F, "nodes_sampled": M, "graph_nodes": int(nodes.shape[0]), "candidates": int(len(cand)), "inserted": k,
            "erode_mesh_3_rounds": {"gpu_ms": t_erode, "cpu_ms": c_erode, "cpu_kind": kind, "cpu_cores": 1},
            "sample_nodes": {"gpu_ms": t_sample, "cpu_ms": c_sample, "cpu_kind": kind, "cpu_cores": 1},
            "nearest_node": {"gpu_ms": t_near, "cpu_ms": c_near, "cpu_kind": "port", "cpu_cores": 1,
                             "gpu_gpairs_per_s": V * nodes.shape[0] / t_near / 1e6},
            "greedy_insert": {"gpu_ms": t_greedy, "cpu_ms": c_greedy, "cpu_kind": "port", "cpu_cores": 1},
        }
    # truncated region: (1) the TSDF of a real integration (c2: sphere at 512^3, source frame), (2) a wavy sheet with a 15 % band
    w = Workload("c2")
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None, device=device)
    vol.integrate({"im": w.frame(0), "id": 0})
    D = 512
    g = torch.Generator(device=device).manual_seed(3)
    z = torch.linspace(-6, 6, D, device=device).view(1, 1, D).expand(D, D, D)
    sheet = (z + 0.3 * torch.randn((D, D, 1), device=device, generator=g)).clamp(-1, 1).contiguous()
    for name, t in (("c2_source_frame_512", vol._tsdf_vol), ("wavy_sheet_512", sheet)):
        reg = torch.empty(t.shape, dtype=torch.uint8, device=device)
        dims = [int(x) for x in t.shape]
        ms = gpu_ms(lambda: _capi.check(lib.ofx_truncated_region(_capi.ptr(t), dims[0], dims[1], dims[2], 1.2, _capi.ptr(reg), S())))
        th = t.cpu().numpy()
        c_ms, want = cpu_ms(lambda: O.truncated_region(th, 1.2), reps=2)
        assert np.array_equal(reg.cpu().numpy().astype(bool), want)
        gbs = 5.0 * t.numel() / ms / 1e6
        out["ops"][f"truncated_region_{name}"] = {
            "gpu_ms": ms, "cpu_ms": c_ms, "cpu_kind": "port", "cpu_cores": cores, "voxels": int(t.numel()),
            "band_frac": float((t.abs() <= 0.9).float().mean()), "region_frac": float(want.mean()),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                         "peak_source": peak_src, "bytes_formula": "5*N (4 read + 1 written)"}}
    # marching cubes inside the truncated region of the same real volume (count -> prefix sum -> emit -> unique -> vertices)
    region = vol.get_truncated_region(1.2)
    mc = lambda: TSDFVolume.marching_cubes(vol._tsdf_vol, region, 0.0, vol._color_vol)      # noqa: E731
    ms = gpu_ms(mc, reps=3)
    v_d, f_d, _, _ = mc()
    th, rh = vol._tsdf_vol.cpu().numpy(), region.cpu().numpy()
    c_ms, (ov_, of_, _) = cpu_ms(lambda: O.marching_cubes(th, rh, 0.0), reps=1)
    assert np.array_equal(v_d.cpu().numpy(), ov_) and np.array_equal(f_d.cpu().numpy(), of_)
    out["ops"]["marching_cubes_c2_source_frame_512"] = {
        "gpu_ms": ms, "cpu_ms": c_ms, "cpu_kind": "port", "cpu_cores": 1, "voxels": int(vol._tsdf_vol.numel()),
        "vertices": int(v_d.shape[0]), "triangles": int(f_d.shape[0]),
        "note": "gpu_ms includes the prefix sum and the key sort (torch.cumsum / torch.unique) and one host read of the triangle count"}
    out["parity"] = "every GPU result above was compared with its CPU baseline's result: identical"
    print(json.dumps(out), flush=True)


def main():
    args = parse()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)                          # anything a library prints (e.g. "NCCL version ...") goes to stderr
    sys.stdout = real_stdout
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        try:        # NCCL's kernels on a high-priority stream: they are tiny and must not queue behind the fused kernel
            opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
            dist.init_process_group("nccl", device_id=device, pg_options=opts)
        except Exception:      # noqa: BLE001
            dist.init_process_group("nccl", device_id=device)
    from occlusionfusion_b200.synthetic import Workload

    if args.graph_ops:
        if rank == 0:
            bench_graph_ops(args, device)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    main_res = bench_workload(args.workload, args, rank, world, device)
    main_res.setdefault("e2e", None)
    also = {}
    if world == 1 and args.also:
        for name in [n for n in args.also.split(",") if n and n != args.workload]:
            r = bench_workload(name, args, rank, world, device, with_e2e=True)
            also[name] = {"value": r["value"], "unit": UNIT, "ms_per_step": r["ms_per_step"],
                          "kernel_ms": r["kernel_ms_local"], "updated_per_frame": r["updated"],
                          "skin_valid_per_frame": r["skin_valid"], "bricks_per_frame": r["bricks"],
                          "voxels": int(np.prod(r["workload"].dims)),
                          "achieved_gbs": r["achieved_gbs"], "e2e": r["e2e"],
                          "config": workload_config(r["workload"], world)}
    cb = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(Workload(args.workload), 6, args.cpu_planes)
    if rank == 0:
        w = main_res["workload"]
        peak, peak_src = hbm_peak()
        line = {
            "metric": METRIC, "value": main_res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": main_res["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (f64 projection/TSDF average)", "data": "synthetic",
            "config": workload_config(w, world, main_res.get("pipelined", False)),
            "voxels": int(np.prod(w.dims)), "updated_per_frame": main_res["updated"],
            "skin_valid_per_frame": main_res["skin_valid"], "bricks_per_frame": main_res["bricks"],
            "voxels_visited_per_s": float(np.prod(w.dims)) / (main_res["ms_per_step"] * 1e-3),
            "wall_ms_per_step": main_res["wall_ms_per_step"],
            "e2e": main_res["e2e"], "gpu_launches": main_res["launches_per_step"] * args.steps,
            "clocks": main_res["clocks"],
            "roofline": {"bound": "hbm", "achieved": main_res["achieved_gbs"], "peak": peak, "unit": "GB/s",
                         "frac": main_res["achieved_gbs"] / peak,
                         "traffic": (ncu_traffic(w.name) or {}).get("dram_bytes_per_launch") if world == 1 else None,
                         "traffic_source": (ncu_traffic(w.name) or {}).get("source") if world == 1 else None,
                         "peak_source": peak_src,
                         "kernel": "fused_warped_persistent<K,VEC,OBS32> (+ brick_cull_kernel: "
                                   "CUDA events around ofx_integrate)",
                         "kernel_ms": main_res["kernel_ms_local"],
                         "algorithmic_bytes_per_launch": main_res["alg_bytes"],
                         "bytes_formula": "24*N_updated(this GPU) + 8*W*H + 60*M"},
            "cpu_baseline": cb, "also": also,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
