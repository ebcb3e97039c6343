"""TEST INFRASTRUCTURE ONLY -- numpy front-end of the CPU oracle (oracle/ofx_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  The product package never does.

`Oracle` mirrors the stages of the reference CPU path one to one (reference file:line in
ofx_oracle.c); `OracleVolume` strings them together the way TSDFVolume.integrate
(tsdf.py:378-494) + WarpField.deform_tsdf (warpfield.py:369-380) do, including the cached
skinning of WarpField.skin_tsdf (warpfield.py:131-141).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libofx_oracle.so")

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("ofx_oracle.c", "ofx_oracle_graph.c", "ofx_oracle_mc.c")]
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(s) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "libofx_oracle.so"])
    return _LIB_PATH


def _load():
    if not os.path.exists(_LIB_PATH):
        build()
    lib = C.CDLL(_LIB_PATH)
    lib.ofx_oracle_expf.restype = C.c_float
    lib.ofx_oracle_expf.argtypes = [C.c_float]
    lib.ofx_oracle_vox2world.argtypes = [_f32p, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
    lib.ofx_oracle_pack_frame.argtypes = [_f32p, C.c_int, C.c_int, _f32p, _f32p]
    lib.ofx_oracle_skin.restype = C.c_int
    lib.ofx_oracle_skin.argtypes = [_f32p, C.c_int64, _f32p, C.c_int, C.c_int, C.c_double,
                                    _i32p, _f32p, _u8p]
    lib.ofx_oracle_warp.argtypes = [_f32p, C.c_int64, _i32p, _f32p, _u8p, C.c_int, _f32p, _f32p, _f32p]
    lib.ofx_oracle_check_visibility.argtypes = [_f32p, C.c_int64, _f32p, C.c_int, C.c_int, _f64p,
                                                C.c_double, _u8p, _f64p]
    lib.ofx_oracle_set_threads.argtypes = [C.c_int]
    lib.ofx_oracle_max_threads.restype = C.c_int
    lib.ofx_oracle_integrate.restype = C.c_int64
    lib.ofx_oracle_integrate.argtypes = [_f32p, _f32p, _f32p, C.c_int64, _f32p, _u8p, _f32p, _f32p,
                                         C.c_int, C.c_int, _f64p, C.c_double, C.c_double]
    lib.ofx_oracle_nearest_node.argtypes = [_f32p, C.c_int64, _f32p, C.c_int, _f32p, _i32p]
    lib.ofx_oracle_greedy_update.restype = C.c_int
    lib.ofx_oracle_greedy_update.argtypes = [_f32p, _i32p, C.c_int, _f32p, C.c_int, C.c_double, _i32p]
    lib.ofx_oracle_sample_nodes.restype = C.c_int
    lib.ofx_oracle_sample_nodes.argtypes = [_f32p, _u8p, C.c_int, C.c_float, C.c_int, _i32p, _f32p]
    lib.ofx_oracle_erode_mesh.argtypes = [_i32p, C.c_int, C.c_int, C.c_int, C.c_int, _u8p]
    lib.ofx_oracle_truncated_region.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_double, _u8p]
    lib.ofx_oracle_mc_triangles.restype = C.c_int64
    lib.ofx_oracle_mc_triangles.argtypes = [_f32p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p]
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _load()
    return _lib


def set_threads(n):
    lib().ofx_oracle_set_threads(int(n))


def max_threads():
    return int(lib().ofx_oracle_max_threads())


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def expf(x):
    return np.array([lib().ofx_oracle_expf(float(v)) for v in np.ravel(x)], np.float32).reshape(np.shape(x))


def vox2world(origin, voxel_size, dims, x0=0):
    """dims = (Dx, Dy, Dz) of the slab starting at global x-plane x0."""
    Dx, Dy, Dz = (int(d) for d in dims)
    out = np.empty((Dx * Dy * Dz, 3), np.float32)
    lib().ofx_oracle_vox2world(_c(origin, np.float32), float(voxel_size), int(x0), Dx, Dy, Dz, out)
    return out


def pack_frame(im):
    im = _c(im, np.float32)
    assert im.ndim == 3 and im.shape[0] == 6
    H, W = im.shape[1:]
    color, depth = np.empty((H, W), np.float32), np.empty((H, W), np.float32)
    lib().ofx_oracle_pack_frame(im, H, W, color, depth)
    return color, depth


def skin(points, nodes, K, coverage):
    points, nodes = _c(points, np.float32), _c(nodes, np.float32)
    n, M = len(points), len(nodes)
    anchors = np.empty((n, K), np.int32)
    weights = np.empty((n, K), np.float32)
    valid = np.empty(n, np.uint8)
    rc = lib().ofx_oracle_skin(points, n, nodes, M, K, float(coverage), anchors, weights, valid)
    if rc:
        raise ValueError(f"oracle skin: bad K={K} for M={M}")
    return anchors, weights, valid.astype(bool)


def warp(points, anchors, weights, valid, R, T):
    points = _c(points, np.float32)
    K = anchors.shape[1]
    out = np.empty_like(points)
    lib().ofx_oracle_warp(points, len(points), _c(anchors, np.int32), _c(weights, np.float32),
                          _c(valid, np.uint8), K, _c(R, np.float32).reshape(-1, 9),
                          _c(T, np.float32), out)
    return out


def check_visibility(points, depth, intr4, trunc):
    points, depth = _c(points, np.float32), _c(depth, np.float32)
    H, W = depth.shape
    valid = np.empty(len(points), np.uint8)
    diff = np.empty(len(points), np.float64)
    lib().ofx_oracle_check_visibility(points, len(points), depth, H, W, _c(intr4, np.float64),
                                      float(trunc), valid, diff)
    return valid.astype(bool), diff


def integrate(tsdf, weight, colour, cam_pts, valid_points, depth, color_im, intr4, trunc, obs_weight=1.0):
    """In place on the three float32 volumes (any shape, C order). Returns #updated voxels."""
    for a in (tsdf, weight, colour):
        assert a.dtype == np.float32 and a.flags.c_contiguous
    depth, color_im = _c(depth, np.float32), _c(color_im, np.float32)
    H, W = depth.shape
    return int(lib().ofx_oracle_integrate(tsdf.reshape(-1), weight.reshape(-1), colour.reshape(-1),
                                          tsdf.size, _c(cam_pts, np.float32), _c(valid_points, np.uint8),
                                          depth, color_im, H, W, _c(intr4, np.float64),
                                          float(trunc), float(obs_weight)))


def nearest_node(points, nodes):
    """warpfield.py:463: float32 distance to / index of the nearest node."""
    points, nodes = _c(points, np.float32).reshape(-1, 3), _c(nodes, np.float32)
    dist, idx = np.empty(len(points), np.float32), np.empty(len(points), np.int32)
    lib().ofx_oracle_nearest_node(points, len(points), nodes, len(nodes), dist, idx)
    return dist, idx


def greedy_update(verts, cand, nodes, coverage):
    """embedded_deformation_graph.py:516-536: accepted vertex ids, in acceptance order."""
    verts, nodes, cand = _c(verts, np.float32), _c(nodes, np.float32).reshape(-1, 3), _c(cand, np.int32)
    acc = np.empty(max(len(cand), 1), np.int32)
    n = lib().ofx_oracle_greedy_update(verts, cand, len(cand), nodes, len(nodes), float(coverage), acc)
    return acc[:n].copy()


def sample_nodes(verts, non_eroded, coverage, use_only_non_eroded=True):
    """graph_proc.cpp:79-136 (randomShuffle = false): (node positions, node vertex ids)."""
    verts = _c(verts, np.float32)
    ne = _c(np.asarray(non_eroded).reshape(-1), np.uint8)
    idx, pos = np.empty(max(len(verts), 1), np.int32), np.empty((max(len(verts), 1), 3), np.float32)
    n = lib().ofx_oracle_sample_nodes(verts, ne, len(verts), float(coverage), int(use_only_non_eroded), idx, pos)
    return pos[:n].copy(), idx[:n].copy()


def erode_mesh(n_verts, faces, n_iter, min_neighbours):
    """graph_proc.cpp:17-77: (V,) bool, True = not eroded."""
    faces = _c(faces, np.int32).reshape(-1, 3)
    out = np.empty(max(int(n_verts), 1), np.uint8)
    lib().ofx_oracle_erode_mesh(faces, len(faces), int(n_verts), int(n_iter), int(min_neighbours), out)
    return out[:int(n_verts)].astype(bool)


def truncated_region(tsdf, max_diff):
    """tsdf.py:706-745: bool volume."""
    tsdf = _c(tsdf, np.float32)
    out = np.empty(tsdf.shape, np.uint8)
    lib().ofx_oracle_truncated_region(tsdf, *tsdf.shape, float(max_diff), out)
    return out.astype(bool)


def mc_vertex_positions(vol, keys, level=0.0):
    """Vertices named by grid-edge keys (3 * lower-corner index + axis) -> (n,3) float32 voxel coordinates: the lower corner
    plus t = (level - v0) / (v1 - v0) along the edge's axis, float32 throughout."""
    vol = _c(vol, np.float32)
    W, H, D = vol.shape
    lower, axis = keys // 3, (keys % 3).astype(np.int64)
    stride = np.array([H * D, D, 1], np.int64)[axis]
    v0, v1 = vol.reshape(-1)[lower], vol.reshape(-1)[lower + stride]
    t = (np.float32(level) - v0) / (v1 - v0)
    pos = np.stack([lower // (H * D), (lower // D) % H, lower % D], 1).astype(np.float32)
    pos[np.arange(len(keys)), axis] += t
    return pos


def marching_cubes(vol, mask=None, level=0.0):
    """Classic marching cubes with the generated case table (ofx_oracle_mc.c): verts (n,3) f32 in voxel coordinates (axis
    order of the array), faces (T,3) i32.  Vertices unique per grid edge, sorted by edge key; faces in cube scan order."""
    vol = _c(vol, np.float32)
    W, H, D = vol.shape
    m = None if mask is None else _c(mask, np.uint8)
    mp = None if m is None else m.ctypes.data
    nt = lib().ofx_oracle_mc_triangles(vol, mp, W, H, D, float(level), None)
    keys = np.empty(3 * nt, np.int64)
    lib().ofx_oracle_mc_triangles(vol, mp, W, H, D, float(level), keys.ctypes.data)
    ukeys, inverse = np.unique(keys, return_inverse=True)
    return mc_vertex_positions(vol, ukeys, level), inverse.reshape(-1, 3).astype(np.int32), ukeys


def origin_translations(R, t, nodes):
    """warpfield.py:407-408: T_i = -(R_i @ g_i) + g_i + t_i, in the estimator's dtypes."""
    return -np.array([R[i] @ nodes[i] for i in range(len(nodes))]) + nodes + t


class OracleVolume:
    """The reference sequence TSDFVolume + WarpField on the CPU, via the oracle stages."""

    def __init__(self, dims, origin, voxel_size, intr4, trunc=0.04, x0=0):
        self.dims = tuple(int(d) for d in dims)
        self.origin = np.asarray(origin, np.float32)
        self.voxel_size = float(voxel_size)
        self.intr4 = np.asarray(intr4, np.float64)
        self.trunc = float(trunc)
        self.tsdf = np.ones(self.dims, np.float32)        # tsdf.py:133
        self.weight = np.zeros(self.dims, np.float32)     # tsdf.py:138
        self.colour = np.zeros(self.dims, np.float32)     # tsdf.py:141
        self.world_pts = vox2world(self.origin, self.voxel_size, self.dims, x0)
        self._skin = None

    def set_graph(self, nodes, coverage, K=None):
        self.nodes = np.asarray(nodes, np.float32)
        self.coverage = float(coverage)
        self.K = min(len(self.nodes), 4) if K is None else int(K)   # warpfield.py:60
        self._skin = None                                           # warpfield.py:493

    def set_transforms(self, R, T):
        self.R, self.T = np.asarray(R, np.float32), np.asarray(T, np.float32)

    def skin_tsdf(self):
        if self._skin is None:
            self._skin = skin(self.world_pts, self.nodes, self.K, self.coverage)
        return self._skin

    def deform_tsdf(self):
        a, w, v = self.skin_tsdf()
        return warp(self.world_pts, a, w, v, self.R, self.T), v

    def integrate(self, im, identity_warp, obs_weight=1.0):
        color_im, depth_im = pack_frame(im)
        if identity_warp:                                           # tsdf.py:395-398
            pts, valid = self.world_pts, np.ones(len(self.world_pts), bool)
        else:
            pts, valid = self.deform_tsdf()                         # tsdf.py:401
        return integrate(self.tsdf, self.weight, self.colour, pts, valid, depth_im, color_im,
                         self.intr4, self.trunc, obs_weight)
