"""Seeded synthetic inputs for the fusion hot path: camera, volume
geometry (through the same formulas as the reference constructor), RGB-D frames, graph nodes
and per-frame node transforms.  Pure numpy; used by tests, bench.py and smoke() for BOTH the
CUDA path and the CPU oracle, so the two always see identical inputs.

Workloads
  "band"   a sphere (r = 0.45 m) in front of the camera, nodes on its camera-facing surface;
           only voxels within 4*node_coverage of the surface are skin-valid (the realistic case).
  "dense"  nodes on a jittered 3-D lattice through the whole volume, depth = a plane at the far
           face: every voxel is skin-valid and every in-frustum voxel updates (the HBM-bound
           case the roofline figure is quoted on).  With surface=True the plane is tilted and cuts
           the volume at 80 % of its depth (a real surface band; voxels behind it do not update).
"""
import types

import numpy as np

CONFIGS = {
    # name: dims, voxel size (m), nodes, K, image (H, W), scene
    "tiny": dict(dim=[40, 48, 64], voxel=0.02, n_nodes=120, K=4, H=120, W=160, scene="band"),
    "c0": dict(dim=128, voxel=0.02, n_nodes=400, K=4, H=480, W=640, scene="band"),
    "c1": dict(dim=256, voxel=0.01, n_nodes=1000, K=4, H=480, W=640, scene="band"),
    "c2": dict(dim=512, voxel=0.005, n_nodes=2000, K=4, H=480, W=640, scene="band"),
    "c4": dict(dim=1024, voxel=0.0025, n_nodes=4000, K=8, H=720, W=1280, scene="band"),
    "c1-dense": dict(dim=256, voxel=0.005, n_nodes=1000, K=4, H=480, W=640, scene="dense"),
    "c2-dense": dict(dim=512, voxel=0.0025, n_nodes=2197, K=4, H=480, W=640, scene="dense"),
    # the same dense volume and graph, but the depth image is a tilted plane THROUGH the volume: voxels in front of it see
    # free space, a band of +-trunc around it takes the float64 running average, voxels behind it are not updated
    "c2-surface": dict(dim=512, voxel=0.0025, n_nodes=2197, K=4, H=480, W=640, scene="dense", surface=True),
}
SEED0 = 20221
# seed offsets are fixed per workload (adding a workload must not reseed the others)
SEED_INDEX = {"c0": 0, "c1": 1, "c1-dense": 2, "c2": 3, "c2-dense": 4, "c4": 5, "tiny": 6, "c2-surface": 7}


def rodrigues(rv):
    """rotation vectors (M,3) -> rotation matrices (M,3,3), float64."""
    th = np.linalg.norm(rv, axis=1)
    k = rv / np.maximum(th, 1e-12)[:, None]
    Kx = np.zeros((len(rv), 3, 3))
    Kx[:, 0, 1], Kx[:, 0, 2] = -k[:, 2], k[:, 1]
    Kx[:, 1, 0], Kx[:, 1, 2] = k[:, 2], -k[:, 0]
    Kx[:, 2, 0], Kx[:, 2, 1] = -k[:, 1], k[:, 0]
    s, c = np.sin(th)[:, None, None], np.cos(th)[:, None, None]
    return np.eye(3)[None] + s * Kx + (1 - c) * (Kx @ Kx)


def _hash_colour(H, W, seed):
    v, u = np.mgrid[0:H, 0:W].astype(np.uint32)
    out = np.empty((3, H, W), np.float32)
    for ch in range(3):
        h = (u * np.uint32(73856093)) ^ (v * np.uint32(19349663)) ^ np.uint32((seed * 83492791 + ch * 2654435761) & 0xFFFFFFFF)
        h ^= h >> np.uint32(13)
        h *= np.uint32(0x5bd1e995)
        h ^= h >> np.uint32(15)
        out[ch] = (h & np.uint32(255)).astype(np.float32) / np.float32(255.0)
    return out


class Workload:
    """One synthetic configuration.  Attributes: name, H, W, intr (fx,fy,cx,cy float64), bbox,
    max_depth, fopt (namespace for the TSDFVolume constructor), dims, voxel_size, origin, nodes
    (M,3) f32, coverage, K, scene."""

    def __init__(self, name="tiny", seed=None, **override):
        cfg = dict(CONFIGS[name])
        cfg.update(override)
        self.name, self.cfg = name, cfg
        self.seed = SEED0 + SEED_INDEX[name] if seed is None else seed
        rng = np.random.default_rng(self.seed)
        H, W = cfg["H"], cfg["W"]
        self.H, self.W = H, W
        f = 525.0 * W / 640
        self.intr = np.array([f, f, (W - 1) / 2, (H - 1) / 2])
        dims = cfg["dim"]
        dims = [dims] * 3 if isinstance(dims, int) else list(dims)
        vs = cfg["voxel"]
        # choose bbox / max_depth so that the reference constructor (tsdf.py:75-107) yields
        # exactly this voxel size: z extent = max_depth = Dz*vs is the largest scaled extent
        Z = dims[2] * vs
        half_w = min(W / 2.0, np.floor(0.5 * dims[0] * vs * f / Z * 0.99))
        half_h = min(H / 2.0, np.floor(0.5 * dims[1] * vs * f / Z * 0.99))
        cx, cy = self.intr[2], self.intr[3]
        self.bbox = (cx - half_w, cy - half_h, cx + half_w, cy + half_h)
        self.max_depth = Z
        self.fopt = types.SimpleNamespace(voxel_dim=dims, source_frame=0, skip_rate=1)
        from .tsdf import TSDFVolume
        K3 = np.eye(3)
        K3[0, 0], K3[1, 1], K3[0, 2], K3[1, 2] = self.intr
        bnds, vdim, vsize = TSDFVolume.volume_geometry(self.bbox, Z, K3, self.fopt)
        self.dims = tuple(int(d) for d in vdim)
        self.voxel_size = float(vsize)
        self.origin = bnds[:, 0].astype(np.float32)
        self.trunc = 0.04
        self.K = cfg["K"]
        self.scene = cfg["scene"]
        M = cfg["n_nodes"]
        ext = np.array(self.dims) * self.voxel_size
        if self.scene == "band":
            self.centre = np.array([0.0, 0.0, 0.5 * Z])
            self.radius = min(0.45, 0.25 * Z)
            # Fibonacci points on the camera-facing hemisphere, jittered into general position
            i = np.arange(M) + 0.5
            zc = -(i / M)                                    # cos(polar) in (-1, 0): facing the camera
            phi = np.pi * (1 + 5 ** 0.5) * i
            rr = np.sqrt(1 - zc * zc)
            d = np.stack([rr * np.cos(phi), rr * np.sin(phi), zc], 1)
            nodes = self.centre + self.radius * d
            self.coverage = float(np.round(np.sqrt(2 * np.pi * self.radius ** 2 / M), 4))
        else:
            n = int(round(M ** (1 / 3)))
            sp = ext / n
            gi = np.stack(np.meshgrid(*(np.arange(n),) * 3, indexing="ij"), -1).reshape(-1, 3)
            nodes = self.origin.astype(np.float64) + (gi + 0.5) * sp
            self.coverage = float(np.round(0.5 * sp.max(), 4))
            self.plane_depth = float(self.origin[2] + ext[2])
            self.surface = bool(cfg.get("surface", False))
            if self.surface:
                self.plane_n = np.array([0.3, 0.2, 1.0]) / np.linalg.norm([0.3, 0.2, 1.0])
                self.plane_c = float(self.plane_n[2] * (self.origin[2] + 0.8 * ext[2]))
        nodes = nodes + rng.uniform(-1e-3, 1e-3, nodes.shape)
        self.nodes = nodes.astype(np.float32)
        self.M = len(self.nodes)

    # ---- frames ---------------------------------------------------------------------------------
    def frame(self, f):
        """(6,H,W) float32: RGB in [0,1] + XYZ metres (layout model/dataset.py:202-206)."""
        H, W = self.H, self.W
        fx, fy, cx, cy = self.intr
        v, u = np.mgrid[0:H, 0:W].astype(np.float64)
        d = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], 0)
        if self.scene == "band":
            c = self.centre + np.array([0.004, -0.003, 0.002]) * f
            a = (d * d).sum(0)
            b = -2 * (d * c[:, None, None]).sum(0)
            cc = float(c @ c) - self.radius ** 2
            disc = b * b - 4 * a * cc
            t = np.where(disc > 0, (-b - np.sqrt(np.maximum(disc, 0))) / (2 * a), 0.0)
        elif getattr(self, "surface", False):
            t = (self.plane_c - 0.001 * f) / (d * self.plane_n[:, None, None]).sum(0)      # ray . n = c: camera z of the plane
        else:
            t = np.full((H, W), self.plane_depth - 0.001 * f)
        im = np.empty((6, H, W), np.float32)
        im[:3] = _hash_colour(H, W, self.seed + f)
        im[3:] = (d * t).astype(np.float32)
        return im

    # ---- node transforms ------------------------------------------------------------------------
    def estimate(self, f, nodes=None):
        """Estimator-style dict for frame f (registration_fusion.py:372-375): smooth rotation and
        translation fields evaluated at the nodes."""
        g = (self.nodes if nodes is None else nodes).astype(np.float64)
        rv = 0.15 * np.sin(2 * np.pi * g / 1.3 + 0.05 * f)
        R = rodrigues(rv).astype(np.float32)
        t = (0.02 * np.cos(2 * np.pi * g / 0.9 + 0.05 * f)).astype(np.float32)
        return dict(node_rotations=R, node_translations=t,
                    deformed_nodes_to_target=(g + t).astype(np.float32),
                    source_frame_id=0, target_frame_id=f)

    def graph(self):
        from .embedded_deformation_graph import EDGraph
        return EDGraph(nodes=self.nodes, graph_generation_parameters={"node_coverage": self.coverage})
