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
    ap.add_argument("--also", default="c2,c1,c2-surface,c3", help="extra workloads reported under 'also' (N=1; at N=8 'c4' is added); '' = none")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA graph replay")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (investigation runs)")
    ap.add_argument("--no-parity", action="store_true", help="skip the post-run parity check against the CPU oracle")
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
    """SM clock and clock-event reasons sampled DURING the timed region: NVML polled in-process every millisecond by a
    thread (the timed region is tens of milliseconds: `nvidia-smi -lms` cannot resolve it).  start()/stop() bracket
    the region; only samples taken between mark_begin() and mark_end() count."""
    REASONS = (("hw_slowdown", "nvmlClocksEventReasonHwSlowdown"), ("hw_thermal_slowdown", "nvmlClocksEventReasonHwThermalSlowdown"),
               ("sw_thermal_slowdown", "nvmlClocksEventReasonSwThermalSlowdown"), ("sw_power_cap", "nvmlClocksEventReasonSwPowerCap"),
               ("hw_power_brake", "nvmlClocksEventReasonHwPowerBrakeSlowdown"))

    def __init__(self, index=0):
        self.rows, self.index, self.th, self.err = [], index, None, None
        self._stop = threading.Event()
        self.t_begin = self.t_end = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # CUDA_VISIBLE_DEVICES may renumber the devices: go through the PCI bus id of the torch device
            import torch
            bus = torch.cuda.get_device_properties(self.index).pci_bus_id if hasattr(torch.cuda.get_device_properties(self.index), "pci_bus_id") else None
            h = None
            if bus is not None:
                for i in range(pynvml.nvmlDeviceGetCount()):
                    hi = pynvml.nvmlDeviceGetHandleByIndex(i)
                    if int(pynvml.nvmlDeviceGetPciInfo(hi).bus) == int(bus):
                        h = hi
                        break
            self.h = h if h is not None else pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:      # noqa: BLE001
            self.err = f"NVML unavailable ({type(e).__name__}: {e})"
            return
        self.th = threading.Thread(target=self._poll, daemon=True)
        self.th.start()

    def _poll(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:       # noqa: BLE001
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append((time.perf_counter(), float(mhz), int(mask)))
            except Exception as e:      # noqa: BLE001
                self.err = str(e)
                return
            time.sleep(0.001)

    def mark_begin(self):
        self.t_begin = time.perf_counter()

    def mark_end(self):
        self.t_end = time.perf_counter()

    def stop(self):
        if self.th is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [self.err or "NVML unavailable"], "samples": 0}
        if self.t_end is None:
            self.mark_end()
        self._stop.set()
        self.th.join(timeout=1)
        inside = [r for r in self.rows if (self.t_begin or 0) <= r[0] <= self.t_end]
        note = None
        if len(inside) < 3:
            inside, note = self.rows, "timed region too short to sample: all samples of the run (warm-up included) reported"
        reasons = set()
        for _, _, mask in inside:
            for name, attr in self.REASONS:
                if mask & int(getattr(self.nv, attr, 0)):
                    reasons.add(name)
        out = {"sm_mhz": float(np.median([r[1] for r in inside])) if inside else None, "sm_max_mhz": self.max_mhz,
               "reasons": sorted(reasons), "samples": len(inside), "source": "NVML polled in-process every ms during the timed region"}
        if note:
            out["note"] = note
        return out


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
        if rank == 0:
            sampler.start()
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
        barrier()
        sampler.mark_begin()
        e0, e1 = mk(), mk()
        t_wall0 = time.perf_counter()
        e0.record(stream)
        if graph is not None:
            graph.replay()
        else:
            run_pipeline()
        e1.record(stream)
        barrier()
        sampler.mark_end()
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
        for i in range(20):                                    # ~10 ms of the same load before the region: the sampler sees the clocks settle
            prep(i)
            graph.replay() if graph is not None else core()
        barrier()
        sampler.mark_begin()
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
        sampler.mark_end()
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
               local_updated=local_upd, launches_per_step=3, graph=graph is not None, pipelined=pipelined)   # frame preparation (+ node transforms), brick cull, fused kernel
    # roofline of the fused kernel on this rank's shard (rank 0 reports its own shard)
    alg_bytes = 24.0 * local_upd + 8.0 * H * W + 60.0 * w.M
    res["alg_bytes"] = alg_bytes
    res["achieved_gbs"] = alg_bytes / (float(kern_ms_local(evs, K)) * 1e-3) / 1e9
    res["kernel_ms_local"] = kern_ms_local(evs, K)

    # ---- e2e: public API, host buffers, H2D + D2H inside the timed region ------------------------
    if with_e2e and not args.no_e2e:
        res["e2e"] = bench_e2e(w, vol, wf, sf, frames, ests, args, rank, world, device)
    # fast-path indicator: the share of observed voxels whose TSDF is exactly 1 (free space seen again keeps 1
    # without the float64 average); the rest takes the float64 running average
    obs = (vol._weight_vol > 0)
    n_obs = int(obs.sum().item())
    n_one = int((obs & (vol._tsdf_vol == 1.0)).sum().item())
    if world > 1:
        cnt2 = torch.tensor([n_obs, n_one], dtype=torch.int64, device=device)
        dist.all_reduce(cnt2, op=dist.ReduceOp.SUM)
        n_obs, n_one = int(cnt2[0]), int(cnt2[1])
    res["free_space_share"] = n_one / max(1, n_obs)
    if not args.no_parity:
        res["parity"] = parity_check(w, vol, sf, rank, world, device)
    del vol, wf, sf, packets, flush
    torch.cuda.empty_cache()
    return res


def kern_ms_local(evs, K):
    return sum(c.elapsed_time(d) for _, _, c, d in evs) / K


def bench_e2e(w, vol, wf, sf, frames, ests, args, rank, world, device):
    """Per step, through the public classes, starting from HOST memory: rank 0 takes the camera frame (page-locked
    numpy) + the estimator dict, WarpField.update_transformations (host numpy), host->device copy of RGB+Z and of R|T,
    [N>1: the packet travels to the other ranks], the kernels, and the frame's counters come back to the host.
    The public calls do not wait for the device: the upload of frame t+1 overlaps the kernels of frame t, and the
    counters a step reads are those of the frame BEFORE it (the last ones are read after the loop, inside the timed
    region)."""
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

    def host_side(i):
        """rank 0: the estimator's output -> origin-form node transforms (WarpField.update_transformations)"""
        est = dict(ests[i % n_frames])
        fid[0] += 1
        est["target_frame_id"] = fid[0]
        wf.frame_id = vol.frame_id = fid[0] - 1
        wf.update_transformations(est)
        vol.frame_id = fid[0]
        return frames[1 + i % n_frames], wf.rotations, wf.translations

    def run(n):
        """n frames; returns the updated voxels summed over them (every frame's counters are read on the host)"""
        upd = 0
        if world == 1:
            vol.frame_id = fid[0]
            for i in range(n):
                est = dict(ests[i % n_frames])
                fid[0] += 1
                est["target_frame_id"] = fid[0]
                wf.update_transformations(est)
                vol.integrate({"im": frames[1 + i % n_frames], "id": fid[0]})
                if i > 0:
                    upd += vol.last_stats(previous=True)["updated"]          # frame i-1: never stalls frame i
            return upd + vol.last_stats()["updated"]
        # N > 1: the pipelined transport on host frames (sharded.py: pipeline_step(next_packet=(im, R, T)))
        sf.pipeline_begin(H, W, host_side(0) if rank == 0 else None)
        for i in range(n):
            nxt = host_side(i + 1) if (rank == 0 and i + 1 < n) else None
            sf.pipeline_step(nxt, last=(i + 1 == n))
            if i >= 2:
                upd += int(sf.counters_host(i - 2)[:, 0].sum())               # gathered two exchanges ago: no stall
        sf.pipeline_end()
        for t in range(max(0, n - 2), n):
            upd += int(sf.counters_host(t)[:, 0].sum())
        return upd

    run(Wm)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    upd = run(K)
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    tmax = torch.tensor([dt], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dt = float(tmax[0])
    h2d = 4 * 4 * H * W + 48 * w.M
    return {"value": upd / dt, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 32 * world,
            "ms_per_step": dt * 1e3 / K, "updated_per_frame": upd / K,
            "timing": "host wall clock around K steps of the public calls on host frames; uploads overlap the kernels, "
                      "every frame's counters are read on the host inside the timed region (one frame late)"}


# ------------------------------------------------------------------------------------------------
def parity_check(w, vol, sf, rank, world, device, n_warped=3, seed=4321):
    """OUTSIDE the timed region: resets this rank's shard, drives the source frame + n_warped warped frames through the
    SAME driver the timed run used (N > 1: the pipelined transport -- symmetric-memory packets, two streams, one collective
    per frame; N = 1: the fused call on device packets), then compares a seeded sample of THIS rank's voxels (uniform,
    around the surface, whole z-rows) with the CPU oracle replaying the same frames through its point operators
    (skin -> warp -> integrate).  Bit-exact or it counts as a mismatch.  Returns (checked, mismatches) summed over ranks."""
    import torch
    import torch.distributed as dist
    from occlusionfusion_b200 import _capi
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    O.build()
    O.set_threads(max(1, min(8, (os.cpu_count() or 8) // max(1, world))))
    H, W = w.H, w.W
    lib = _capi.load()
    _capi.check(lib.ofx_volume_reset(vol._h, _capi.stream_ptr(device)))
    frames = [w.frame(i) for i in range(n_warped + 1)]           # every rank can render the synthetic frames itself
    ests = [w.estimate(i) for i in range(1, n_warped + 1)]
    Ts = [(-np.einsum("nij,nj->ni", e["node_rotations"], w.nodes) + w.nodes + e["node_translations"]).astype(np.float32) for e in ests]
    # source frame (identity warp), sequential driver
    if rank == 0:
        sf.fill_packet(frames[0], np.tile(np.eye(3, dtype=np.float32), (w.M, 1, 1)), np.zeros((w.M, 3), np.float32))
    sf.step(H, W, identity_warp=True)
    packets = []
    for i in range(n_warped):
        if rank == 0:
            sf.fill_packet(frames[i + 1], ests[i]["node_rotations"], Ts[i])
        torch.cuda.synchronize(device)
        packets.append(sf.buf.clone())
    if world > 1:
        sf.pipeline_begin(H, W, packets[0] if rank == 0 else None)
        for i in range(n_warped):
            sf.pipeline_step(packets[i + 1] if (rank == 0 and i + 1 < n_warped) else None, last=(i + 1 == n_warped))
        sf.pipeline_end()
    else:
        for pk in packets:
            sf.buf.copy_(pk)
            v = sf.views
            sf.be.integrate_warped(v["R"], v["T"], v["depth"], v["colour"], 1.0)
    torch.cuda.synchronize(device)

    # the sample: local (lx, iy, iz) of this rank's planes
    Dx, Dy, Dz = w.dims
    nx = len(sf.planes)
    rng = np.random.default_rng(seed + rank)
    n_uni = 60_000
    ijk = [np.stack([rng.integers(0, nx, n_uni), rng.integers(0, Dy, n_uni), rng.integers(0, Dz, n_uni)], 1)]
    if w.scene == "band":
        d = rng.normal(size=(100_000, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        d[:, 2] = -np.abs(d[:, 2])
        p = w.centre + (w.radius + rng.uniform(-0.08, 0.08, (len(d), 1))) * d
        q = np.rint((p - w.origin.astype(np.float64)) / w.voxel_size).astype(np.int64)
        q = q[((q >= 0) & (q < np.array(w.dims))).all(1)]
        lx_of = -np.ones(Dx, np.int64)
        lx_of[sf.planes] = np.arange(nx)
        keep = lx_of[q[:, 0]] >= 0
        ijk.append(np.stack([lx_of[q[keep, 0]], q[keep, 1], q[keep, 2]], 1))
    rows = np.stack([rng.integers(0, nx, 24), rng.integers(0, Dy, 24)], 1)
    ijk.append(np.concatenate([np.concatenate([np.repeat(r[None], Dz, 0), np.arange(Dz)[:, None]], 1) for r in rows]))
    ijk = np.concatenate(ijk)
    flat = np.unique((ijk[:, 0] * Dy + ijk[:, 1]) * Dz + ijk[:, 2])
    lx, iy, iz = flat // (Dy * Dz), (flat // Dz) % Dy, flat % Dz
    px = O.vox2world(w.origin, w.voxel_size, (Dx, 1, 1))[:, 0][sf.planes]
    py = O.vox2world(w.origin, w.voxel_size, (1, Dy, 1))[:, 1]
    pz = O.vox2world(w.origin, w.voxel_size, (1, 1, Dz))[:, 2]
    pts = np.stack([px[lx], py[iy], pz[iz]], 1).astype(np.float32)
    a, wts, valid = O.skin(pts, w.nodes, w.K, w.coverage)
    t = np.ones(len(flat), np.float32)
    wt = np.zeros(len(flat), np.float32)
    c = np.zeros(len(flat), np.float32)
    for f in range(n_warped + 1):
        color_im, depth_im = O.pack_frame(frames[f])
        if f == 0:
            cam_pts, ok = pts, np.ones(len(pts), bool)
        else:
            cam_pts, ok = O.warp(pts, a, wts, valid, ests[f - 1]["node_rotations"], Ts[f - 1]), valid
        O.integrate(t, wt, c, cam_pts, ok, depth_im, color_im, w.intr, w.trunc, 1.0)
    ft = torch.from_numpy(flat).to(device)
    g_t = vol._tsdf_vol.reshape(-1)[ft].cpu().numpy()
    g_w = vol._weight_vol.reshape(-1)[ft].cpu().numpy()
    g_c = vol._color_vol.reshape(-1)[ft].cpu().numpy()
    bad = int(((g_t != t) | (g_w != wt) | (g_c != c)).sum())
    res = torch.tensor([len(flat), bad, int((wt > 0).sum())], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(res, op=dist.ReduceOp.SUM)
    return {"checked": int(res[0]), "mismatches": int(res[1]), "sampled_voxels_updated": int(res[2]), "frames": n_warped + 1,
            "how": "after the timed region every rank resets its shard, reruns source + warped frames through the driver the "
                   "timed run used and compares a seeded voxel sample of its own planes with the CPU oracle, bit for bit"}


def bench_c3(args, device, n_frames=100):
    """BASELINE config 3: a 100-frame sequence at 256^3 with a per-frame graph-node update and node insertion every 10th
    frame, through the drop-in classes the way fusion_with_occlusion/fusion.py:202-221 drives them (the canonical mesh
    for the insertion comes from the volume itself: truncated region -> marching cubes -> erosion -> unreachable vertices ->
    greedy insertion -> edge rebuild -> re-skin).  Host wall clock around synchronised phases; the estimator that would
    produce the transforms is outside the path and outside the timing."""
    import torch
    from occlusionfusion_b200 import EDGraph, TSDFVolume, WarpField
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("c1")
    vol = TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None, device=device)
    frames = [torch.from_numpy(w.frame(f)).pin_memory().numpy() for f in range(8)]

    def timed(fn):
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize(device)
        return (time.perf_counter() - t0) * 1e3, r

    t_src, _ = timed(lambda: vol.integrate({"im": frames[0], "id": 0}))
    graph = EDGraph(tsdf=vol, nodes=np.zeros((0, 3), np.float32), graph_generation_parameters={"node_coverage": w.coverage})
    t_build, _ = timed(lambda: graph.create_graph_from_mesh(*vol.get_canonical_model()[:2]))
    wf = WarpField(graph, vol, None)
    wf.graph_neighbours = w.K
    vol.warpfield = wf
    t_graph, growth, upd_total = [], [int(graph.nodes.shape[0])], 0
    ests = {}
    t_all0 = time.perf_counter()
    t_est = 0.0
    t_frames = 0.0
    for f in range(1, n_frames):
        if f % 10 == 0:
            vol.clear()
            ms, _ = timed(lambda: wf.update_graph())
            t_graph.append(ms)
            growth.append(int(wf.graph.nodes.shape[0]))
        t0 = time.perf_counter()
        est = w.estimate(f, wf.graph.nodes)                     # stands in for the estimator (not part of the path)
        t_est += time.perf_counter() - t0
        t0 = time.perf_counter()
        wf.update_transformations(est)
        vol.integrate({"im": frames[f % len(frames)], "id": f})
        if f > 1 and f % 10 != 1:
            upd_total += vol.last_stats(previous=True)["updated"]
        t_frames += time.perf_counter() - t0
        if f % 10 == 9 or f == n_frames - 1:
            t0 = time.perf_counter()
            upd_total += vol.last_stats()["updated"]            # drains the frame loop before the graph update
            t_frames += time.perf_counter() - t0
    torch.cuda.synchronize(device)
    wall = time.perf_counter() - t_all0 - t_est
    n = n_frames - 1
    return {"workload": f"c3 = {w.name}: {w.dims[0]}^3 voxels @ {w.voxel_size * 1e3:.0f} mm, {w.W}x{w.H}, K={w.K}, {n_frames} frames, "
                        f"update_transformations + integrate every frame from host frames, update_graph every 10th frame",
            "frames": n_frames, "ms_per_frame_all_in": wall * 1e3 / n, "value": upd_total / wall, "unit": UNIT,
            "frame_loop_ms_per_frame": t_frames * 1e3 / n, "update_graph_ms": [round(x, 2) for x in t_graph],
            "update_graph_ms_median": float(np.median(t_graph)) if t_graph else None, "nodes": growth,
            "source_frame_ms": t_src, "graph_from_source_mesh_ms": t_build, "updated_voxels": int(upd_total),
            "timing": "host wall clock; frame loop asynchronous (counters read one frame late), each update_graph synchronised"}


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
        # the same on the device (csrc/ofx_graph_edges.cu): mesh uploaded once, one CTA per node search; wall clock of the
        # whole call (it returns host arrays), compared with the host builder's result
        from occlusionfusion_b200.graph_proc import DeviceMesh, compute_edges_geodesic_d
        mesh_d = DeviceMesh(verts, faces, np.ones(V, bool), device)
        de, dw, dd = -np.ones((M, 8), np.int32), np.zeros((M, 8), np.float32), np.zeros((M, 8), np.float32)
        compute_edges_geodesic_d(mesh_d, nidx, 8, cov, de, dw, dd, True, True)          # warm (pool allocations)
        t_geo_dev = []
        for _ in range(3):
            de[:] = -1; dw[:] = 0; dd[:] = 0
            torch.cuda.synchronize(device)
            t0 = time.perf_counter()
            replayed = compute_edges_geodesic_d(mesh_d, nidx, 8, cov, de, dw, dd, True, True)
            t_geo_dev.append((time.perf_counter() - t0) * 1e3)
        assert np.array_equal(de, ge) and np.array_equal(dw, gw) and np.array_equal(dd, gd)
        out["ops"][f"mesh_{n}x{n}"] = {
            "edges_geodesic_host": {"ms": t_geo, "host_threads": os.cpu_count(), "cpu_ms": c_geo, "cpu_kind": "reference",
                                    "cpu_cores": 1, "where": "host C++ in libofx (the reference's is host C++ too)"},
            "edges_geodesic_device": {"gpu_ms": float(np.median(t_geo_dev)), "nodes_replayed_for_ties": int(replayed),
                                      "timing": "host wall clock of the whole call (results back on the host), mesh already uploaded"},
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
    # marching cubes inside the truncated region of the same real volume (count -> device scans -> vertices -> faces, one library call)
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
        "note": "one ofx_mc_mesh call (count, two device scans, vertices, faces: own kernels, no sort) + the one host read of the two counts"}
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
    peak, peak_src = hbm_peak()

    def also_entry(r):
        e = {"value": r["value"], "unit": UNIT, "ms_per_step": r["ms_per_step"], "kernel_ms": r["kernel_ms_local"],
             "updated_per_frame": r["updated"], "skin_valid_per_frame": r["skin_valid"], "bricks_per_frame": r["bricks"],
             "voxels": int(np.prod(r["workload"].dims)), "achieved_gbs": r["achieved_gbs"], "e2e": r.get("e2e"),
             "free_space_share_of_observed_voxels": r.get("free_space_share"), "parity": r.get("parity"),
             "roofline": {"bound": "hbm", "achieved": r["achieved_gbs"], "peak": peak, "unit": "GB/s", "frac": r["achieved_gbs"] / peak,
                          "algorithmic_bytes_per_launch": r["alg_bytes"], "kernel_ms": r["kernel_ms_local"], "peak_source": peak_src},
             "config": workload_config(r["workload"], world, r.get("pipelined", False))}
        return e

    names = [n for n in args.also.split(",") if n and n != args.workload] if args.also else []
    def settle():
        """Between workloads (never inside a timed region): the previous workload's volume, warp field and graph reference one
        another, so only the cyclic collector frees them -- at a moment of its choosing, e.g. in the middle of the next
        workload's update_graph, where the cudaFree of gigabytes of volume and skin cache showed up as single 20-70 ms
        updates.  Collect now."""
        import gc
        gc.collect()
        torch.cuda.synchronize(device)

    if world == 1:
        for name in names:
            settle()
            if name == "c3":
                also["c3"] = bench_c3(args, device)
            else:
                also[name] = also_entry(bench_workload(name, args, rank, world, device, with_e2e=True))
    elif world == 8 and args.also:
        # BASELINE config 5 (1024^3, 1280x720, ~4k nodes, K = 8) is quoted on 8 GPUs
        r = bench_workload("c4", args, rank, world, device, with_e2e=True)
        if rank == 0:
            also["c4"] = also_entry(r)
    cb = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(Workload(args.workload), 6, args.cpu_planes)
    if rank == 0:
        w = main_res["workload"]
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
                         "kernel": "fused_warped_persistent<K,OBS1> (+ frame_prep_kernel + brick_cull_kernel: "
                                   "CUDA events around ofx_integrate_warped)",
                         "kernel_ms": main_res["kernel_ms_local"],
                         "algorithmic_bytes_per_launch": main_res["alg_bytes"],
                         "bytes_formula": "24*N_updated(this GPU) + 8*W*H + 60*M"},
            "cpu_baseline": cb, "also": also,
            "parity": main_res.get("parity"),
            "free_space_share_of_observed_voxels": main_res.get("free_space_share"),
        }
        print(json.dumps(line), flush=True)
    bad = (main_res.get("parity") or {}).get("mismatches", 0)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if bad:
        raise SystemExit(f"bench.py: parity check FAILED: {bad} sampled voxels differ from the oracle")


if __name__ == "__main__":
    main()
