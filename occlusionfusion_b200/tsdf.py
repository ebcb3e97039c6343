"""TSDFVolume -- drop-in for the reference class of the same name
(/root/reference/fusion_with_occlusion/tsdf.py:37-875) on the non-rigid fusion hot path.

Same constructor and method signatures; the three volumes live in HBM as torch tensors and
every per-voxel computation runs in libofx.so (include/ofx.h).  What the reference keeps as
full-volume host arrays (`world_pts`, `vox_coords`, cached anchors/weights) is never
materialised on the per-frame path; `world_pts` is produced on request.

Surface extraction (tsdf.py:704-809) runs on the device too: truncated region + classic marching cubes + colour unpack
(`get_mesh`, `get_canonical_model`).  Not rebuilt: everything that needs a visualiser.
"""
import ctypes as C
import logging
import os
import pickle

import numpy as np
import torch

from . import _capi

# output capacities of marching_cubes() remembered per (volume shape, device): (vertices, triangles)
_MC_CAPACITY = {}


class TSDFVolume:
    """Volumetric TSDF fusion of RGB-D frames under a non-rigid warp field (tsdf.py:37)."""

    def __init__(self, bbox, max_depth, cam_intr, fopt, visualizer=None, *, device=None,
                 shard=0, n_shards=1, x_block=None):
        """
        Args (as tsdf.py:41-49):
            bbox: (w_min, h_min, w_max, h_max) of the captured region in pixels
            max_depth: maximum depth in the sequence (metres)
            cam_intr: (fx, fy, cx, cy)
            fopt: options namespace; reads voxel_dim or voxel_size, source_frame, skip_rate and,
                  when present, datadir (results folders as tsdf.py:314-319)
            visualizer: kept for signature parity; unused (headless)
        Extra keyword-only arguments (multi-GPU): this process owns shard `shard` of `n_shards`;
        x-planes are dealt in blocks of `x_block` (default Dx/n_shards = contiguous slabs).
        """
        _capi.require_cuda()
        self._lib = _capi.load()
        self.fopt = fopt
        self.vis = visualizer
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)

        self.cam_intr = np.eye(3)                                            # tsdf.py:55-59
        self.cam_intr[0, 0], self.cam_intr[1, 1] = cam_intr[0], cam_intr[1]
        self.cam_intr[0, 2], self.cam_intr[1, 2] = cam_intr[2], cam_intr[3]

        self._vol_bnds, self._vol_dim, self._voxel_size = self.volume_geometry(bbox, max_depth, self.cam_intr, fopt)
        self._color_const = 256 * 256
        self._trunc_margin = 0.04                                            # tsdf.py:127
        self._vol_bnds[:, 1] = self._vol_bnds[:, 0] + self._vol_dim * self._voxel_size
        self._vol_origin = self._vol_bnds[:, 0].copy(order="C").astype(np.float32)   # tsdf.py:129

        Dx, Dy, Dz = (int(d) for d in self._vol_dim)
        self.n_shards, self.shard = int(n_shards), int(shard)
        if Dx % self.n_shards:
            raise ValueError(f"voxel_dim[0]={Dx} is not divisible by n_shards={n_shards}")
        self.x_block = Dx // self.n_shards if x_block is None else int(x_block)
        self._nx_local = Dx // self.n_shards
        self._local_dim = (self._nx_local, Dy, Dz)

        with torch.cuda.device(self.device):
            self._tsdf_vol = torch.empty(self._local_dim, dtype=torch.float32, device=self.device)
            self._weight_vol = torch.empty_like(self._tsdf_vol)
            self._color_vol = torch.empty_like(self._tsdf_vol)
            desc = _capi.VolumeDesc()
            desc.dims[:] = [Dx, Dy, Dz]
            desc.n_shards, desc.shard, desc.x_block = self.n_shards, self.shard, self.x_block
            desc.origin[:] = [float(x) for x in self._vol_origin]
            desc.voxel_size, desc.trunc_margin = float(self._voxel_size), float(self._trunc_margin)
            desc.tsdf, desc.weight, desc.colour = (self._tsdf_vol.data_ptr(), self._weight_vol.data_ptr(),
                                                   self._color_vol.data_ptr())
            h = _capi.C.c_void_p()
            _capi.check(self._lib.ofx_volume_create(_capi.C.byref(h), _capi.C.byref(desc)))
            self._h = h
            _capi.check(self._lib.ofx_volume_reset(self._h, _capi.stream_ptr(self.device)))  # tsdf=1, w=0, c=0
            self._stats = torch.zeros(_capi.OFX_N_STATS, dtype=torch.int64, device=self.device)
        self._intr4 = _capi.intr4(self.cam_intr)
        # Two per-frame sets (raw frame, packed colour, depth, counters) used in turn, an upload stream and events:
        # the host->device copy of frame t+1 runs on the upload stream while the kernels of frame t run on the
        # caller's stream; nothing in integrate() waits for the device (counters are read lazily, last_stats()).
        self._set = 0
        self._fd_views = [None, None]
        self._comp = None            # the compute stream of the integrate() call in progress (update() reuses it)
        self._frame_host = [None, None]     # pinned staging for frames that arrive in pageable memory
        self._frame_dev = [None, None]
        self._color_ims = [None, None]
        self._depth_ims = [None, None]
        self._stats_sets = [self._stats, torch.zeros_like(self._stats)]
        self._stats_host = [torch.zeros(_capi.OFX_N_STATS, dtype=torch.int64).pin_memory() for _ in range(2)]
        self._h2d_done = [torch.cuda.Event(), torch.cuda.Event()]
        self._set_done = [torch.cuda.Event(), torch.cuda.Event()]     # the kernels that read set j have finished
        self._set_used = [False, False]
        self._stats_ready = [None, None]                               # event after the counters' device->host copy
        with torch.cuda.device(self.device):
            self._copy_stream = torch.cuda.Stream(device=self.device)
        self._color_im_dev = self._depth_im_dev = None
        self._world_pts = None

        if hasattr(fopt, "datadir") and fopt.datadir is not None:            # tsdf.py:314-319
            self.savepath = os.path.join(fopt.datadir, "results")
            for sub in ("", "tsdf", "visible_nodes", "canonical_model"):
                os.makedirs(os.path.join(self.savepath, sub), exist_ok=True)
        else:
            self.savepath = None
        self.log = logging.getLogger(__name__)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and getattr(self, "_lib", None) is not None:
            self._lib.ofx_volume_destroy(h)
            self._h = None

    # ------------------------------------------------------------------ geometry (host, O(1))
    @staticmethod
    def volume_geometry(bbox, max_depth, cam_intr3x3, fopt):
        """Volume bounds, dims and voxel size exactly as tsdf.py:75-118 derives them."""
        w_min, h_min, w_max, h_max = bbox
        K = cam_intr3x3
        zs = np.array([0, max_depth, max_depth, max_depth, max_depth])
        view_frust_pts = np.array([
            (np.array([0, w_min, w_min, w_max, w_max]) - K[0, 2]) * zs / K[0, 0],
            (np.array([0, h_min, h_max, h_min, h_max]) - K[1, 2]) * zs / K[1, 1],
            zs,
        ])
        vol_bnds = np.asarray([np.min(view_frust_pts, axis=1), np.max(view_frust_pts, axis=1)]).T
        assert vol_bnds.shape == (3, 2), "[!] `vol_bnds` should be of shape (3, 2)."
        if hasattr(fopt, "voxel_dim"):                                       # tsdf.py:95-107
            voxel_dim = fopt.voxel_dim
            if type(voxel_dim) == int:
                vol_dim = np.array([voxel_dim, voxel_dim, voxel_dim])
            elif type(voxel_dim) == list:
                assert len(voxel_dim) == 3, f"Voxel dimension should have length = 3 found {voxel_dim}"
                vol_dim = np.array(voxel_dim)
            else:
                voxel_dim = np.asarray(voxel_dim)
                assert voxel_dim.shape[0] == 3, f"Voxel dimension should have length = 3 found {voxel_dim}"
                vol_dim = voxel_dim
            voxel_size = ((vol_bnds[:, 1] - vol_bnds[:, 0]) / vol_dim).max()
        elif hasattr(fopt, "voxel_size"):                                    # tsdf.py:110-118
            voxel_size = float(fopt.voxel_size)
            vol_dim = np.ceil((vol_bnds[:, 1] - vol_bnds[:, 0]) / voxel_size).copy(order="C").astype(int)
        else:
            raise AttributeError("fopt needs voxel_dim or voxel_size")
        return vol_bnds, vol_dim, voxel_size

    @property
    def world_pts(self):
        """(N,3) float32 voxel positions (tsdf.py:307), materialised on first access only."""
        if self._world_pts is None:
            n = int(self._lib.ofx_volume_num_voxels(self._h))
            out = torch.empty((n, 3), dtype=torch.float32, device=self.device)
            _capi.check(self._lib.ofx_volume_world_points(self._h, _capi.ptr(out), _capi.stream_ptr(self.device)))
            self._world_pts = out.cpu().numpy()
        return self._world_pts

    # ------------------------------------------------------------------ per-frame path
    def update(self, im, frame_id):
        """Update RGB-D frame info (tsdf.py:545-572): (6,H,W) float32, RGB in [0,1] + XYZ.

        Only RGB + Z (4 of the 6 channels) are uploaded -- on a separate upload stream, into the per-frame set that the
        frame before last used, so the copy overlaps the previous frame's kernels -- and folded into the packed colour
        image and the depth image on the device (same float32 op order).  A frame in page-locked memory is DMA'd
        straight from the caller's buffer: do not overwrite it before wait_uploaded() returns; a frame in pageable
        memory is copied into a pinned staging buffer first and may be reused at once."""
        im_t = im if isinstance(im, torch.Tensor) else None
        shape = tuple(im.shape)
        assert len(shape) == 3 and shape[0] >= 4, \
            f"Input not correct\nDepth:{shape[1:]} should be (H,W)\nColor:{shape} should be (H,W,3)"
        if shape[0] != 6:
            raise ValueError("expected a (6,H,W) RGB+XYZ frame")
        H, W = shape[1:]
        j = self._set = 1 - self._set
        if self._frame_dev[j] is None or tuple(self._frame_dev[j].shape[1:]) != (H, W):
            self._frame_dev[j] = torch.empty((4, H, W), dtype=torch.float32, device=self.device)
            self._color_ims[j] = torch.empty((H, W), dtype=torch.float32, device=self.device)
            self._depth_ims[j] = torch.empty((H, W), dtype=torch.float32, device=self.device)
            self._frame_host[j] = None
            self._fd_views[j] = (self._frame_dev[j][:3], self._frame_dev[j][3])
        comp = self._comp if self._comp is not None else torch.cuda.current_stream(self.device)
        fd = self._frame_dev[j]
        if im_t is not None and im_t.is_cuda:
            fd[:3].copy_(im_t[:3])
            fd[3].copy_(im_t[-1])
        else:
            src = im_t if im_t is not None else torch.from_numpy(np.asarray(im, dtype=np.float32))
            if not src.is_pinned():
                if self._frame_host[j] is None:
                    self._frame_host[j] = torch.empty((4, H, W), dtype=torch.float32, pin_memory=True)
                elif self._set_used[j]:
                    self._h2d_done[j].synchronize()       # the DMA of two frames ago has left this staging buffer
                self._frame_host[j][:3].copy_(src[:3])
                self._frame_host[j][3].copy_(src[-1])
                src4 = self._frame_host[j]
            else:
                src4 = None
            cs = self._copy_stream
            if self._set_used[j]:
                cs.wait_event(self._set_done[j])          # the kernels of two frames ago are done with this set
            # (set_stream both ways instead of `with torch.cuda.stream(cs)`: the context manager looks the current stream
            # up twice, ~20 us of Python per frame; `comp` is the current stream, known here)
            torch.cuda.set_stream(cs)
            try:
                if src4 is not None:
                    fd.copy_(src4, non_blocking=True)
                else:
                    fd3, fdz = self._fd_views[j]
                    fd3.copy_(src[:3], non_blocking=True)
                    fdz.copy_(src[-1], non_blocking=True)
                self._h2d_done[j].record(cs)
            finally:
                torch.cuda.set_stream(comp)
            comp.wait_event(self._h2d_done[j])
        self.h2d_bytes_last_frame = fd.numel() * 4
        self._color_im_dev, self._depth_im_dev, self._stats = self._color_ims[j], self._depth_ims[j], self._stats_sets[j]
        _capi.check(self._lib.ofx_pack_frame(_capi.ptr(fd), 4, H, W, _capi.ptr(self._color_im_dev),
                                             _capi.ptr(self._depth_im_dev), _capi.stream_ptr(self.device)))
        self.im = im
        if hasattr(self, "frame_id"):                                        # tsdf.py:569-570
            assert self.frame_id + self.fopt.skip_rate == frame_id, \
                f"Updating image data failed. Previous frame:{self.frame_id}, current frame:{frame_id} not integrated."
        self.frame_id = frame_id

    def wait_uploaded(self):
        """Blocks until the most recent frame has left the caller's (page-locked) buffer."""
        self._h2d_done[self._set].synchronize()

    @property
    def color_im(self):
        return self._color_im_dev.cpu().numpy()

    @property
    def depth_im(self):
        return self._depth_im_dev.cpu().numpy()

    def integrate(self, image_data, obs_weight=1.):
        """Integrate an RGB-D frame into the TSDF volume (tsdf.py:378-494).

        image_data: {"im": (6,H,W) float32, "id": int}.  Source frame -> identity warp
        (tsdf.py:395-398); any other frame is warped by self.warpfield (tsdf.py:401).
        Returns once the work is enqueued; last_stats() (or any read of the volume) waits for it."""
        with torch.cuda.device(self.device):
            comp = self._comp = torch.cuda.current_stream(self.device)      # looked up once per frame (8 us of Python a call)
            try:
                self.update(image_data["im"], image_data["id"])
            finally:
                self._comp = None
            identity = self.frame_id == self.fopt.source_frame
            H, W = self._depth_im_dev.shape
            S = _capi.stream_ptr(self.device)
            done = False
            if not identity:
                wf = self.warpfield
                assert wf.frame_id == self.frame_id, \
                    f"Warpfield maps to:{wf.frame_id}th frame but TSDF maps to:{self.frame_id}th frame"  # warpfield.py:374
                wf._sync_graph()
                if wf._transforms_dirty:
                    # new node transforms ride in the frame-preparation launch (ofx_integrate_warped)
                    R_dev, T_dev = wf._stage_transforms()
                    _capi.check(self._lib.ofx_integrate_warped(self._h, _capi.ptr(R_dev), _capi.ptr(T_dev),
                                                               _capi.ptr(self._depth_im_dev), _capi.ptr(self._color_im_dev),
                                                               H, W, self._intr4, float(obs_weight), _capi.ptr(self._stats), S))
                    wf._transforms_dirty = False
                    done = True
            if not done:
                _capi.check(self._lib.ofx_integrate(self._h, _capi.ptr(self._depth_im_dev), _capi.ptr(self._color_im_dev),
                                                    H, W, self._intr4, float(obs_weight), int(identity),
                                                    _capi.ptr(self._stats), S))
            # the frame's counters go to page-locked host memory behind the kernels; nobody waits here
            j = self._set
            self._stats_host[j].copy_(self._stats, non_blocking=True)
            self._set_done[j].record(comp)
            self._set_used[j] = True
            self._stats_ready[j] = self._set_done[j]

    def last_stats(self, previous=False):
        """Counters of the last integrate(): dict(updated, skin_valid, bricks, weight_sum); waits for that frame.
        previous=True: the counters of the frame BEFORE the last one (None when there is none) -- in a frame loop
        that read never stalls the device, since the last frame is still free to run."""
        j = 1 - self._set if previous else self._set
        ev = self._stats_ready[j]
        if ev is None:
            if previous:
                return None
            s = self._stats.cpu().tolist()
        else:
            ev.synchronize()
            s = self._stats_host[j].tolist()
        return dict(updated=s[_capi.STAT_UPDATED], skin_valid=s[_capi.STAT_SKIN_VALID],
                    bricks=s[_capi.STAT_BRICKS], weight_sum=s[_capi.STAT_WEIGHT_SUM])

    # ------------------------------------------------------------------ callers either side
    def check_visibility(self, points):
        """For each point check that it is not behind the image depth (tsdf.py:599-612).
        Returns (valid (n,) bool, depth_diff (n,) float64)."""
        pts = _capi.to_device(points, torch.float32, self.device).reshape(-1, 3)
        n = pts.shape[0]
        valid = torch.empty(n, dtype=torch.uint8, device=self.device)
        diff = torch.empty(n, dtype=torch.float64, device=self.device)
        H, W = self._depth_im_dev.shape
        with torch.cuda.device(self.device):
            _capi.check(self._lib.ofx_check_visibility(_capi.ptr(pts), n, _capi.ptr(self._depth_im_dev), H, W,
                                                       self._intr4, float(self._trunc_margin), _capi.ptr(valid),
                                                       _capi.ptr(diff), _capi.stream_ptr(self.device)))
        return valid.cpu().numpy().astype(bool), diff.cpu().numpy()

    def get_visible_nodes(self):
        """Visibility of the deformed graph nodes in the current frame (tsdf.py:614-638)."""
        assert self.warpfield.frame_id == self.frame_id
        deformed_nodes = self.warpfield.get_deformed_nodes()
        visible_nodes, _ = self.check_visibility(deformed_nodes)
        if self.savepath is not None:
            np.save(os.path.join(self.savepath, "visible_nodes", f"{self.frame_id}.npy"), visible_nodes)
        return visible_nodes

    def get_volume(self):
        """(tsdf, colour, weight) as host float32 arrays of this shard (tsdf.py:673-680)."""
        return self._tsdf_vol.cpu().numpy(), self._color_vol.cpu().numpy(), self._weight_vol.cpu().numpy()

    def save_volume(self, datapath):                                         # tsdf.py:682-687
        data = [np.expand_dims(d, axis=0) for d in self.get_volume()]
        with open(datapath, "wb") as f:
            pickle.dump(np.concatenate(data, axis=0), f)

    def load_volume(self, datapath):                                         # tsdf.py:689-702
        with open(datapath, "rb") as f:
            data = pickle.load(f)
        self._tsdf_vol.copy_(torch.from_numpy(np.ascontiguousarray(data[0], dtype=np.float32)))
        self._color_vol.copy_(torch.from_numpy(np.ascontiguousarray(data[1], dtype=np.float32)))
        self._weight_vol.copy_(torch.from_numpy(np.ascontiguousarray(data[2], dtype=np.float32)))

    @staticmethod
    def compute_truncated_region(tsdf_vol, max_diff):
        """The mask the reference hands to marching cubes (tsdf.py:706-745, call at :791): True where |tsdf| <= 0.9, the
        voxel is not on the volume boundary and no neighbour differs by more than max_diff.  Runs on the device
        (ofx_truncated_region).  tsdf_vol: (Dx,Dy,Dz) float32 numpy array -> bool numpy array (as the reference), or a
        CUDA tensor -> CUDA bool tensor (no copies)."""
        _capi.require_cuda()
        as_numpy = not isinstance(tsdf_vol, torch.Tensor)
        dev = tsdf_vol.device if not as_numpy and tsdf_vol.is_cuda else torch.device("cuda", torch.cuda.current_device())
        t = _capi.to_device(tsdf_vol, torch.float32, dev)
        assert t.dim() == 3
        out = torch.empty(t.shape, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _capi.check(_capi.load().ofx_truncated_region(_capi.ptr(t), t.shape[0], t.shape[1], t.shape[2], float(max_diff),
                                                          _capi.ptr(out), _capi.stream_ptr(dev)))
        return out.cpu().numpy().astype(bool) if as_numpy else out.bool()

    def get_truncated_region(self, max_diff=1.2):
        """compute_truncated_region of this volume's TSDF, on the device, as a CUDA bool tensor (tsdf.py:791 uses 1.2)."""
        assert self.n_shards == 1, "the truncated region looks at x-neighbours: gather the shards first (ShardedFusion.gather_volume)"
        return self.compute_truncated_region(self._tsdf_vol, max_diff)

    @staticmethod
    def marching_cubes(tsdf_vol, mask=None, level=0.0, color_vol=None):
        """Classic marching cubes on the device (ofx_mc_mesh: one library call; the scikit-image call of tsdf.py:794 --
        Lewiner's variant there, see DESIGN.md for what differs).  tsdf_vol / mask / color_vol: (Dx,Dy,Dz) numpy arrays or
        CUDA tensors.  Returns CUDA tensors: verts (n,3) f32 in voxel coordinates (array axis order), faces (T,3) i32,
        normals (n,3) f32 (unit gradient, towards increasing values), rgb (n,3) u8 or None.  Vertices are unique per grid
        edge and sorted by edge; faces come in cube order.  One host read (the two counts) after the kernels; the output
        buffers are sized from the previous call on a volume of the same shape (retried with room when they were short)."""
        _capi.require_cuda()
        lib = _capi.load()
        dev = tsdf_vol.device if isinstance(tsdf_vol, torch.Tensor) and tsdf_vol.is_cuda \
            else torch.device("cuda", torch.cuda.current_device())
        t = _capi.to_device(tsdf_vol, torch.float32, dev)
        assert t.dim() == 3
        Dx, Dy, Dz = (int(d) for d in t.shape)
        if mask is None:
            m = None
        elif isinstance(mask, torch.Tensor):
            m = mask.to(dev).to(torch.uint8).contiguous()
        else:
            m = _capi.to_device(np.asarray(mask).astype(np.uint8), torch.uint8, dev)
        c = None if color_vol is None else _capi.to_device(color_vol, torch.float32, dev)
        key = (Dx, Dy, Dz, dev.index)
        cap_v, cap_t = _MC_CAPACITY.get(key, (1 << 14, 1 << 15))
        with torch.cuda.device(dev):
            S = _capi.stream_ptr(dev)
            counts = torch.empty(2, dtype=torch.int32, device=dev)
            while True:
                verts = torch.empty((cap_v, 3), dtype=torch.float32, device=dev)
                normals = torch.empty((cap_v, 3), dtype=torch.float32, device=dev)
                rgb = torch.empty((cap_v, 3), dtype=torch.uint8, device=dev) if c is not None else None
                faces = torch.empty((cap_t, 3), dtype=torch.int32, device=dev)
                _capi.check(lib.ofx_mc_mesh(_capi.ptr(t), _capi.ptr(m), _capi.ptr(c), Dx, Dy, Dz, C.c_float(level), cap_v, cap_t,
                                            _capi.ptr(verts), _capi.ptr(normals), _capi.ptr(rgb), _capi.ptr(faces),
                                            _capi.ptr(counts), S))
                n, T = (int(x) for x in counts.tolist())                         # the one host read
                if n <= cap_v and T <= cap_t:
                    break
                cap_v, cap_t = max(cap_v, n + n // 4 + 64), max(cap_t, T + T // 4 + 64)
            _MC_CAPACITY[key] = (max(cap_v, n + n // 4 + 64), max(cap_t, T + T // 4 + 64))
        return verts[:n], faces[:T], normals[:n], (rgb[:n] if rgb is not None else None)

    def get_mesh(self):
        """(verts, faces, norms, colors) of the zero level set inside the truncated region (tsdf.py:779-809): verts in world
        coordinates, colors uint8 (r, g, b) of the voxel nearest to each vertex."""
        assert self.n_shards == 1, "marching cubes needs the whole volume: gather the shards first (ShardedFusion.gather_volume)"
        region = self.get_truncated_region(1.2)                                # tsdf.py:791
        verts, faces, norms, rgb = self.marching_cubes(self._tsdf_vol, region, 0.0, self._color_vol)
        verts = verts.cpu().numpy() * np.float32(self._voxel_size) + np.asarray(self._vol_origin, dtype=np.float32)   # :798
        return verts, faces.cpu().numpy(), norms.cpu().numpy(), rgb.cpu().numpy()

    def get_canonical_model(self):                                            # tsdf.py:811-821
        if hasattr(self, "canonical_model"):
            return self.canonical_model
        self.canonical_model = self.get_mesh()
        return self.canonical_model

    def get_deformed_model(self):
        """The canonical mesh warped to the current frame (tsdf.py:830-849): (vertices, faces, normals, colors)."""
        if hasattr(self, "deformed_model"):
            return self.deformed_model
        if self.frame_id != self.fopt.source_frame:
            vertices, faces, normals, colors = self.get_canonical_model()
            deformed = self.warpfield.deform_mesh(vertices, normals)
            self.deformed_model = (deformed[0], faces, deformed[1], colors)
        else:
            self.deformed_model = self.get_canonical_model()
        return self.deformed_model

    def clear(self):
        """Drops the per-frame caches so that the next frame recomputes them (tsdf.py:857-877)."""
        for name in ("reduced_graph_dict", "deformed_model", "canonical_model"):
            if hasattr(self, name):
                delattr(self, name)
