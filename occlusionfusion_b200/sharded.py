"""Multi-GPU fusion: the volume is split into x-plane shards, one process per GPU
(`torch.distributed`, NCCL over NVLink/NVSwitch).  Voxels are independent given the frame
(SURVEY.md section 8e): no halo, no data-path collective.  Per frame there is exactly

    1 broadcast   of the frame packet  [R (M,9) | T (M,3) | packed colour (H,W) | depth (H,W)]
                  = 48*M + 8*H*W bytes, from rank 0 (the rank that receives the camera frame and
                  the estimator's node transforms)
    1 all-gather  of OFX_N_STATS int64 counters per shard ("slab statistics")

and, only when the graph changes (node insertion), one more broadcast of the node table.

Shards are blocks of `x_block` planes dealt round-robin (default 4 = one brick): the view
frustum is a pyramid inside its bounding box, so contiguous slabs would leave the outer GPUs
nearly idle; dealt blocks give every GPU the same share of in-frustum bricks.

The reference has no multi-device path at all; this module is new.
"""
import os

import numpy as np
import torch
import torch.distributed as dist

N_STATS = 4


def plane_map(Dx, n_shards, shard, x_block):
    """Global x index of each local plane of `shard` (same formula as include/ofx.h)."""
    assert Dx % (x_block * n_shards) == 0, "Dx must be a multiple of x_block * n_shards"
    lx = np.arange(Dx // n_shards)
    return (lx // x_block) * x_block * n_shards + shard * x_block + lx % x_block


def default_x_block(Dx, n_shards, brick_x=4):
    """Smallest brick-aligned block that divides the volume evenly; falls back to contiguous slabs."""
    if Dx % (brick_x * n_shards) == 0:
        return brick_x
    return Dx // n_shards


class FramePacket:
    """Layout of the one buffer broadcast per frame (float32 words)."""

    def __init__(self, M, H, W):
        self.M, self.H, self.W = int(M), int(H), int(W)
        self.o_R = 0
        self.o_T = self.o_R + 9 * self.M
        self.o_colour = -(-(self.o_T + 3 * self.M) // 4) * 4          # 16-byte aligned images
        self.o_depth = self.o_colour + self.H * self.W
        self.numel = -(-(self.o_depth + self.H * self.W) // 4) * 4    # whole 16-byte words (peer copy)

    @property
    def nbytes(self):
        return 4 * self.numel

    def alloc(self, device):
        return torch.zeros(self.numel, dtype=torch.float32, device=device)

    def views(self, buf):
        M, H, W = self.M, self.H, self.W
        return dict(R=buf[self.o_R:self.o_R + 9 * M].view(M, 3, 3), T=buf[self.o_T:self.o_T + 3 * M].view(M, 3),
                    colour=buf[self.o_colour:self.o_colour + H * W].view(H, W),
                    depth=buf[self.o_depth:self.o_depth + H * W].view(H, W))


class ShardedFusion:
    """Drives one shard per rank.  `shard_backend` owns the shard's volume and implements
        set_graph(nodes (M,3) tensor, coverage, K), set_transforms(R, T tensors),
        integrate(depth, colour tensors, identity_warp, obs_weight) -> stats tensor int64[N_STATS],
        pack_frame(im) -> (colour, depth) tensors            [rank 0 only]
        get_volume() -> (tsdf, colour, weight) numpy arrays of the local planes
    The CUDA backend is `CudaShard` below; CPU tests inject an oracle-backed stand-in."""

    def __init__(self, shard_backend, dims, rank, world, x_block, device, group=None):
        self.be, self.dims = shard_backend, tuple(int(d) for d in dims)
        self.rank, self.world, self.x_block = int(rank), int(world), int(x_block)
        self.device, self.group = device, group
        self.planes = plane_map(self.dims[0], self.world, self.rank, self.x_block)
        self.packet = self.buf = None
        self.M = 0

    # -- graph (rare) ----------------------------------------------------------------------------
    def set_graph(self, nodes, coverage, K):
        """rank 0 passes the node table (others may pass None); every rank ends up with it."""
        meta = torch.zeros(3, dtype=torch.float64, device=self.device)
        if self.rank == 0:
            meta[0], meta[1], meta[2] = len(nodes), float(coverage), int(K)
        if self.world > 1:
            dist.broadcast(meta, src=0, group=self.group)
        M, coverage, K = int(meta[0].item()), float(meta[1].item()), int(meta[2].item())
        nd = torch.empty((M, 3), dtype=torch.float32, device=self.device)
        if self.rank == 0:
            nd.copy_(torch.as_tensor(np.ascontiguousarray(nodes, dtype=np.float32)))
        if self.world > 1:
            dist.broadcast(nd, src=0, group=self.group)
        self.M, self.coverage, self.K = M, coverage, K
        self.be.set_graph(nd, coverage, K)
        self.packet = None
        return nd

    # -- per frame -------------------------------------------------------------------------------
    def _ensure_packet(self, H, W):
        if self.packet is None or (self.packet.H, self.packet.W, self.packet.M) != (H, W, self.M):
            self.packet = FramePacket(self.M, H, W)
            self.buf = self.packet.alloc(self.device)
            self.views = self.packet.views(self.buf)

    def fill_packet(self, im, R, T):
        """rank 0: stage the frame (packed on the device) and the node transforms into the packet."""
        H, W = im.shape[-2:]
        self._ensure_packet(H, W)
        colour, depth = self.be.pack_frame(im)
        self.views["colour"].copy_(colour)
        self.views["depth"].copy_(depth)
        self.views["R"].copy_(torch.as_tensor(np.ascontiguousarray(R, dtype=np.float32)), non_blocking=True)
        self.views["T"].copy_(torch.as_tensor(np.ascontiguousarray(T, dtype=np.float32)), non_blocking=True)

    def step(self, H, W, identity_warp=False, obs_weight=1.0, gather_stats=True):
        """Broadcast the packet rank 0 filled, integrate the local shard, all-gather the counters.
        Returns an int64 tensor (world, N_STATS) (or the local (N_STATS,) when gather_stats=False)."""
        self._ensure_packet(H, W)
        if self.world > 1:
            dist.broadcast(self.buf, src=0, group=self.group)
        v = self.views
        stats = self._integrate(v, identity_warp, obs_weight)
        if not gather_stats:
            return stats
        if self.world == 1:
            return stats.view(1, N_STATS)
        out = torch.empty(self.world * N_STATS, dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(out, stats.contiguous().view(-1), group=self.group)
        return out.view(self.world, N_STATS)

    def _integrate(self, v, identity_warp, obs_weight, stats=None):
        """node transforms + integrate of the packet views `v`: one library call where the backend has it
        (ofx_integrate_warped: the transforms ride in the frame-preparation launch)."""
        kw = {} if stats is None else {"stats": stats}
        if identity_warp:
            return self.be.integrate(v["depth"], v["colour"], True, obs_weight, **kw)
        fused = getattr(self.be, "integrate_warped", None)
        if fused is not None:
            return fused(v["R"], v["T"], v["depth"], v["colour"], obs_weight, **kw)
        self.be.set_transforms(v["R"], v["T"])
        return self.be.integrate(v["depth"], v["colour"], False, obs_weight, **kw)

    # -- per frame, pipelined ----------------------------------------------------------------------
    # Two packets, two streams (SURVEY.md 8e "scaling risk"): while the shard integrates frame t the
    # next frame's packet is already being broadcast, and the counters of frame t are gathered behind
    # its kernel.  Order on the communication stream:
    #     bcast(t+1) | wait kernel(t) | all-gather(t) | bcast(t+2) | ...
    # so bcast(t+2) overwrites packet t%2 only after kernel(t) has read it, and overlaps kernel(t+1).
    # Rank 0 hands over frame t+1 (a device-resident packet, e.g. what fill_packet staged) when it
    # submits frame t: one frame of look-ahead.
    def pipeline_setup(self, H, W):
        """Allocates the second packet, the communication stream and the events (once; not capturable)."""
        self._ensure_packet(H, W)
        dev = self.device
        if getattr(self, "_pbuf", None) is not None and self._pbuf[0] is self.buf:
            return
        # Packet transport.  Preferred: the two packets live in symmetric memory (every rank maps every
        # rank's buffer through NVLink / NVSwitch); rank 0 fills its copy, one tiny collective tells the
        # others it is there, and each rank PULLS it with libofx's peer-copy kernel -- 7 concurrent reads of
        # 2.6 MB instead of NCCL's ring broadcast (measured at 8 GPUs: the ring took longer than the
        # per-GPU kernel).  Fallback when symmetric memory is unavailable: dist.broadcast.
        self._peer_src = None
        if self.world > 1 and self.device.type == "cuda" and not os.environ.get("OFX_NO_SYMM"):
            try:
                import torch.distributed._symmetric_memory as symm
                grp = self.group if self.group is not None else dist.group.WORLD
                bufs = [symm.empty(self.packet.numel, dtype=torch.float32, device=dev) for _ in range(2)]
                hdls = [symm.rendezvous(b, grp) for b in bufs]
                for b in bufs:
                    b.zero_()
                self._peer_src = [h.get_buffer(0, (self.packet.numel,), torch.float32) for h in hdls]
                self._symm = (bufs, hdls)                       # keep the mappings alive
                self.buf = bufs[0]
                self.views = self.packet.views(self.buf)
                self._pbuf = bufs
            except Exception as e:                              # noqa: BLE001
                import sys
                print(f"[sharded] symmetric memory unavailable ({type(e).__name__}: {e}); using dist.broadcast", file=sys.stderr)
                self._peer_src = None
        # Large packets (BASELINE config 5: 7.6 MB) are PUSHED once through the NVSwitch multicast mapping of the symmetric
        # buffers instead of being pulled seven times out of rank 0 (its NVLink egress is the bound there): rank 0 stages the
        # packet in a local buffer, and after the per-frame all-gather (= every rank is done with the frame that used this
        # packet buffer) copies it to the multicast address (ofx_multicast_copy); a barrier on the symmetric-memory signal
        # pads tells the peers it has landed.  OFX_MULTICAST=1 / 0 forces it on / off; default: packets of 4 MB and more.
        self._mc = None
        if self._peer_src is not None:
            mc = [int(getattr(h, "multicast_ptr", 0) or 0) for h in self._symm[1]]
            want = os.environ.get("OFX_MULTICAST")
            if all(mc) and (want == "1" or (want is None and self.packet.nbytes >= (4 << 20))):
                self._mc = mc
                if self.rank == 0 and os.environ.get("OFX_DEBUG"):
                    import sys
                    print(f"[sharded] packets of {self.packet.nbytes} bytes go out through the NVSwitch multicast mapping", file=sys.stderr)
                self._pstage = [self.packet.alloc(dev) if self.rank == 0 else None for _ in range(2)]
        if self._peer_src is None:
            self._pbuf = [self.buf, self.packet.alloc(dev)]
        self._pviews = [self.packet.views(b) for b in self._pbuf]
        # where rank 0 stages a packet before it is published: the packet buffer itself, or the local buffer of the multicast push
        self._sbuf = self._pstage if (self._mc and self.rank == 0) else self._pbuf
        self._sviews = [self.packet.views(b) for b in self._sbuf] if (self._mc and self.rank == 0) else self._pviews
        self._pstats = [torch.zeros(N_STATS, dtype=torch.int64, device=dev) for _ in range(2)]
        self._pgath = [torch.zeros(self.world * N_STATS, dtype=torch.int64, device=dev) for _ in range(2)]
        # gathered counters also land in page-locked host memory behind each exchange: read lazily (counters_host)
        self._pgath_host = [torch.zeros(self.world * N_STATS, dtype=torch.int64).pin_memory() if dev.type == "cuda"
                            else torch.zeros(self.world * N_STATS, dtype=torch.int64) for _ in range(2)]
        self._ev_gath = [torch.cuda.Event() for _ in range(2)] if dev.type == "cuda" else [None, None]
        # host-frame staging on the rank that receives the camera frames (pipeline_step(next_packet=(im, R, T)))
        self._hraw = [None, None]
        self._hviews = [None, None]
        self._hRT = [None, None]
        self._hfr = [None, None]
        self._ev_staged = [None, None]
        self._comm = torch.cuda.Stream(device=dev, priority=-1)   # its few small operations go first
        self._ev_bcast = [torch.cuda.Event() for _ in range(2)]
        self._ev_done = [torch.cuda.Event() for _ in range(2)]
        self._HW = (H, W)
        # Look-ahead on the communication stream while frame t integrates (OFX_LOOKAHEAD): 1 = frame t+1's node transforms
        # and frustum cull (ofx_prepare_frame), 2 = everything of frame t+1 but the fused kernel (ofx_stage_frame: frame
        # image, depth tiles, node transforms, cull with the depth-range test).  The staging kernels only get SM slots
        # in the tail of the persistent kernel (it owns every slot; leaving one slot per SM free costs it 15-20 %,
        # tools/shard_probe.py), so what is hidden is their launch latency and part of their run time.
        la = os.environ.get("OFX_LOOKAHEAD", "")
        self._lookahead = int(la) if la.isdigit() and hasattr(self.be, "prepare") else 0
        if self._lookahead == 2 and not hasattr(self.be, "stage"):
            self._lookahead = 1

    def pipeline_begin(self, H, W, first_packet=None):
        self.pipeline_setup(H, W)
        dev = self.device
        self._comm.wait_stream(torch.cuda.current_stream(dev))
        self._pipe_t = 0
        for st in self._pstats:
            st.zero_()
        self._comm.wait_stream(torch.cuda.current_stream(dev))
        self._exchange(0, first_packet, prev=None)

    def _exchange(self, t, packet, prev, publish=True, comp=None):
        """Communication-stream work of one step: publish packet t and gather the counters of frame
        `prev` (None: nothing to gather yet).  ONE collective does both jobs: the all-gather of the
        counters is also the hand-shake of the packet transport --
          * it completes on a rank only after rank 0 has entered it, i.e. after rank 0's copy of the new
            packet into its symmetric buffer (stream order on rank 0): the pull below sees the new data;
          * rank 0 leaves it, and may later overwrite the other packet, only after every rank has entered
            it, i.e. after that rank's previous pull (stream order on that rank).
        The pull is a copy-engine transfer from rank 0's mapped buffer: the persistent fused kernel holds
        every SM slot, an SM copy kernel (ofx_peer_copy, OFX_PULL_KERNEL=1) would queue behind it."""
        j = t % 2
        # (set_stream both ways instead of the `with torch.cuda.stream` context: it looks the current stream up twice, ~20 us
        # of Python per frame on the rank that paces the end-to-end loop)
        if comp is None:
            comp = torch.cuda.current_stream(self.device)
        torch.cuda.set_stream(self._comm)
        try:
            if prev is not None:
                self._comm.wait_event(self._ev_done[prev % 2])
            if publish and self.rank == 0 and isinstance(packet, (tuple, list)):
                self._stage_host(j, *packet)                     # HOST frame + node transforms -> packet j (asynchronous)
            elif publish and self.rank == 0 and packet is not None and packet.data_ptr() != self._sbuf[j].data_ptr():
                self._sbuf[j].copy_(packet, non_blocking=True)
            pj = 0 if prev is None else prev % 2
            if self.world > 1:
                dist.all_gather_into_tensor(self._pgath[pj], self._pstats[pj], group=self.group)
            else:
                self._pgath[pj].copy_(self._pstats[pj])
            if prev is not None and self._ev_gath[pj] is not None:
                self._pgath_host[pj].copy_(self._pgath[pj], non_blocking=True)
                self._ev_gath[pj].record(self._comm)
            if self.world > 1 and publish and self._mc:
                from . import _capi
                if self.rank == 0:
                    _capi.check(_capi.load().ofx_multicast_copy(self._mc[j], _capi.ptr(self._sbuf[j]), self.packet.nbytes, 64,
                                                                _capi.stream_ptr(self.device)))
                self._symm[1][j].barrier(0, 10000)          # the push has landed everywhere (10 s time-out: no silent hang)
            elif self.world > 1 and publish:
                if self._peer_src is None:
                    dist.broadcast(self._pbuf[j], src=0, group=self.group)
                elif self.rank != 0:
                    if os.environ.get("OFX_PULL_KERNEL"):
                        from . import _capi
                        _capi.check(_capi.load().ofx_peer_copy(_capi.ptr(self._pbuf[j]), _capi.ptr(self._peer_src[j]),
                                                               self.packet.nbytes, 32, _capi.stream_ptr(self.device)))
                    else:
                        self._pbuf[j].copy_(self._peer_src[j], non_blocking=True)
            if publish and self._lookahead:
                # frame t's node-transform table and frustum cull, still on the communication stream: they
                # overlap the kernel of frame t-1 (their CTAs fill the SM slots its tail frees)
                v = self._pviews[j]
                if self._lookahead == 2:
                    self.be.stage(v["R"], v["T"], v["depth"], v["colour"])
                else:
                    self.be.prepare(v["R"], v["T"], self._HW[0], self._HW[1])
            self._ev_bcast[j].record(self._comm)
        finally:
            torch.cuda.set_stream(comp)
        return self._pgath[pj].view(self.world, N_STATS)

    def _stage_host(self, j, im, R, T):
        """Rank 0, on the communication stream: a HOST frame (6,H,W) float32 (RGB in [0,1] + XYZ, tsdf.py:545-572) and the
        node transforms R (M,3,3), T (M,3) into packet j -- RGB + Z are DMA'd (straight from the caller's buffer when it is
        page-locked: do not overwrite it before wait_staged() of that packet returns), folded into the packet's colour / depth
        images on the device (ofx_pack_frame), R | T go through a pinned staging buffer.  Nothing waits for the device."""
        from . import _capi
        H, W = self._HW
        M = self.M
        dev = self.device
        if self._hraw[j] is None:
            self._hraw[j] = torch.empty((4, H, W), dtype=torch.float32, device=dev)
            self._hRT[j] = torch.empty(12 * M, dtype=torch.float32).pin_memory()
            # views made once: this function runs per frame on the rank that paces the end-to-end loop
            self._hviews[j] = (self._hraw[j][:3], self._hraw[j][3], self._hRT[j].numpy(), self._sbuf[j][:12 * M])
        elif self._ev_staged[j] is not None:
            self._ev_staged[j].synchronize()                     # the staging of two frames ago has left the pinned buffers
        src = im if isinstance(im, torch.Tensor) else torch.from_numpy(np.asarray(im, dtype=np.float32))
        if not src.is_pinned():
            if self._hfr[j] is None:
                self._hfr[j] = torch.empty((4, H, W), dtype=torch.float32).pin_memory()
            self._hfr[j][:3].copy_(src[:3])
            self._hfr[j][3].copy_(src[-1])
            self._hraw[j].copy_(self._hfr[j], non_blocking=True)
        else:
            self._hviews[j][0].copy_(src[:3], non_blocking=True)
            self._hviews[j][1].copy_(src[-1], non_blocking=True)
        v = self._sviews[j]
        _capi.check(_capi.load().ofx_pack_frame(_capi.ptr(self._hraw[j]), 4, H, W, _capi.ptr(v["colour"]), _capi.ptr(v["depth"]),
                                                _capi.stream_ptr(dev)))
        h = self._hviews[j][2]
        h[:9 * M] = np.asarray(R, dtype=np.float32).reshape(-1)
        h[9 * M:] = np.asarray(T, dtype=np.float32).reshape(-1)
        self._hviews[j][3].copy_(self._hRT[j], non_blocking=True)          # R at word 0, T at word 9 M (FramePacket)
        if self._ev_staged[j] is None:
            self._ev_staged[j] = torch.cuda.Event()
        self._ev_staged[j].record(self._comm)

    def wait_staged(self, t):
        """Blocks until the HOST frame and node transforms handed over for frame t (pipeline_begin: t = 0; pipeline_step
        number i: t = i + 1) have been read: a page-locked frame buffer may be overwritten afterwards (a pageable one is
        copied before pipeline_step returns and may be reused at once)."""
        ev = self._ev_staged[t % 2]
        if ev is not None:
            ev.synchronize()

    def counters_host(self, t):
        """The gathered (world, N_STATS) counters of frame t as a host array; waits only for that frame's exchange
        (frame t's counters are gathered by the pipeline_step / pipeline_end call after the one that integrated it)."""
        pj = t % 2
        self._ev_gath[pj].synchronize()
        return self._pgath_host[pj].view(self.world, N_STATS).numpy().copy()

    def pipeline_step(self, next_packet=None, last=False, obs_weight=1.0):
        """Integrate frame t (its packet was published by the previous call) while frame t+1's packet
        (`next_packet` on rank 0: a device-resident packet, or a HOST frame with its node transforms as the tuple
        (im (6,H,W), R (M,3,3), T (M,3)); ignored when `last`) is published and the counters of frame t-1 are
        gathered.  Returns the (world, N_STATS) counter tensor of frame t-1 (meaningless for t = 0); it is valid
        once the communication stream has run.  pipeline_end() returns the last frame's counters."""
        t, j = self._pipe_t, self._pipe_t % 2
        comp = torch.cuda.current_stream(self.device)
        comp.wait_event(self._ev_bcast[j])
        v = self._pviews[j]
        if self._lookahead:
            self.be.integrate(v["depth"], v["colour"], False, obs_weight, stats=self._pstats[j])
        else:
            self._integrate(v, False, obs_weight, stats=self._pstats[j])
        self._ev_done[j].record(comp)
        # issued after integrate(t): the look-ahead stages frame t+1 in the frame set integrate(t) is not using
        out = self._exchange(t + 1, next_packet, prev=t - 1 if t > 0 else None, publish=not last, comp=comp)
        self._pipe_t += 1
        return out

    def pipeline_end(self):
        """Gathers the counters of the frames still in flight and joins the streams; returns the
        (world, N_STATS) counters of the last frame."""
        t = self._pipe_t - 1
        out = None
        if t >= 0:
            out = self._exchange(t + 1, None, prev=t, publish=False)     # gather only
        torch.cuda.current_stream(self.device).wait_stream(self._comm)
        return out

    # -- output (not per frame) ------------------------------------------------------------------
    def gather_volume(self):
        """Full (Dx,Dy,Dz) volumes on rank 0 (None elsewhere): undoes the dealt-block plane mapping."""
        local = [np.ascontiguousarray(a) for a in self.be.get_volume()]
        if self.world == 1:
            full = []
            for a in local:
                f = np.empty(self.dims, np.float32)
                f[self.planes] = a
                full.append(f)
            return tuple(full)
        gathered = [None] * self.world if self.rank == 0 else None
        dist.gather_object((self.planes, local), gathered, dst=0, group=self.group)
        if self.rank != 0:
            return None
        full = [np.empty(self.dims, np.float32) for _ in range(3)]
        for planes, arrs in gathered:
            for f, a in zip(full, arrs):
                f[planes] = a
        return tuple(full)


class CudaShard:
    """shard backend over a TSDFVolume handle: raw C-ABI calls on device tensors (no host hops)."""

    def __init__(self, tsdf_volume):
        from . import _capi
        self._capi, self.vol = _capi, tsdf_volume
        self.lib, self.dev = _capi.load(), tsdf_volume.device
        self.stats = torch.zeros(N_STATS, dtype=torch.int64, device=self.dev)

    def _s(self):
        return self._capi.stream_ptr(self.dev)

    def set_graph(self, nodes, coverage, K):
        c = self._capi
        c.check(self.lib.ofx_set_graph(self.vol._h, c.ptr(nodes.contiguous()), nodes.shape[0], float(coverage), int(K), self._s()))

    def set_transforms(self, R, T):
        c = self._capi
        c.check(self.lib.ofx_set_transforms(self.vol._h, c.ptr(R), c.ptr(T), self._s()))

    def prepare(self, R, T, H, W):
        """Look-ahead: the next frame's node transforms + frustum cull into the other frame set, on the current
        stream (ofx_prepare_frame); the next integrate() then launches only the fused kernel."""
        c = self._capi
        c.check(self.lib.ofx_prepare_frame(self.vol._h, c.ptr(R), c.ptr(T), int(H), int(W), self.vol._intr4, self._s()))

    def stage(self, R, T, depth, colour):
        """Look-ahead, full: everything of the next frame except the fused kernel (ofx_stage_frame) into the other frame
        set, on the current stream; the next integrate() on the same images launches the fused kernel only."""
        c = self._capi
        H, W = depth.shape
        c.check(self.lib.ofx_stage_frame(self.vol._h, c.ptr(R), c.ptr(T), c.ptr(depth), c.ptr(colour), int(H), int(W),
                                         self.vol._intr4, self._s()))

    def pack_frame(self, im):
        self.vol.update(im, getattr(self.vol, "frame_id", -self.vol.fopt.skip_rate) + self.vol.fopt.skip_rate)
        return self.vol._color_im_dev, self.vol._depth_im_dev

    def integrate(self, depth, colour, identity_warp, obs_weight, stats=None):
        c = self._capi
        H, W = depth.shape
        st = self.stats if stats is None else stats
        c.check(self.lib.ofx_integrate(self.vol._h, c.ptr(depth), c.ptr(colour), H, W, self.vol._intr4, float(obs_weight),
                                       int(identity_warp), c.ptr(st), self._s()))
        return st

    def integrate_warped(self, R, T, depth, colour, obs_weight, stats=None):
        """set_transforms + integrate of a warped frame as one library call (ofx_integrate_warped)."""
        c = self._capi
        H, W = depth.shape
        st = self.stats if stats is None else stats
        c.check(self.lib.ofx_integrate_warped(self.vol._h, c.ptr(R), c.ptr(T), c.ptr(depth), c.ptr(colour), H, W, self.vol._intr4,
                                              float(obs_weight), c.ptr(st), self._s()))
        return st

    def get_volume(self):
        return self.vol.get_volume()
