"""ctypes binding of libofx.so (include/ofx.h).  Thin on purpose: torch tensors in, raw device
pointers + sizes + the current CUDA stream out.  There is no CPU fallback: if the library is
missing or no CUDA device is present, every entry point raises.
"""
import ctypes as C
import os

import numpy as np
import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
# OFX_LIB: an experimental build of the same library (tools/ab_bench.sh); never a different implementation
LIB_PATH = os.environ.get("OFX_LIB") or os.path.join(_PKG, "lib", "libofx.so")

OFX_MAX_K = 8
OFX_N_STATS = 4
STAT_UPDATED, STAT_SKIN_VALID, STAT_BRICKS, STAT_WEIGHT_SUM = range(4)

# every symbol include/ofx.h declares (tests check the .so exports all of them)
SYMBOLS = [
    "ofx_abi_version", "ofx_last_error", "ofx_volume_create", "ofx_volume_destroy", "ofx_volume_reset",
    "ofx_volume_num_voxels", "ofx_set_graph", "ofx_set_transforms", "ofx_integrate", "ofx_integrate_warped", "ofx_prepare_frame", "ofx_stage_frame",
    "ofx_volume_world_points", "ofx_volume_skin", "ofx_volume_deform", "ofx_pack_frame", "ofx_skin",
    "ofx_warp_points", "ofx_check_visibility", "ofx_peer_copy", "ofx_multicast_copy",
    "ofx_nearest_node", "ofx_pixel_anchors_euclidean", "ofx_argmin_norm", "ofx_greedy_insert", "ofx_sample_nodes", "ofx_erode_mesh", "ofx_truncated_region",
    "ofx_mc_mesh",
    "ofx_compute_edges_geodesic", "ofx_compute_edges_geodesic_batch", "ofx_edge_weights_host", "ofx_node_and_edge_clean_up", "ofx_compute_clusters", "ofx_compute_clusters_noalloc", "ofx_upsample_index",
    "ofx_compute_edges_geodesic_host", "ofx_node_and_edge_clean_up_host", "ofx_compute_clusters_host", "ofx_downsample_nodes_host",
]


class OfxError(RuntimeError):
    pass


class VolumeDesc(C.Structure):
    _fields_ = [
        ("dims", C.c_int32 * 3), ("n_shards", C.c_int32), ("shard", C.c_int32), ("x_block", C.c_int32),
        ("origin", C.c_float * 3), ("voxel_size", C.c_double), ("trunc_margin", C.c_double),
        ("tsdf", C.c_void_p), ("weight", C.c_void_p), ("colour", C.c_void_p),
    ]


_lib = None


def load():
    """Loads libofx.so; raises OfxError when it has not been built (no fallback path)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OfxError(f"{LIB_PATH} not found: build it with `python -m occlusionfusion_b200.build` "
                       "(there is no CPU or PyTorch fallback for this path)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    lib.ofx_abi_version.restype = C.c_int
    lib.ofx_last_error.restype = C.c_char_p
    lib.ofx_volume_create.argtypes = [C.POINTER(vp), C.POINTER(VolumeDesc)]
    lib.ofx_volume_destroy.argtypes = [vp]
    lib.ofx_volume_reset.argtypes = [vp, vp]
    lib.ofx_volume_num_voxels.argtypes = [vp]
    lib.ofx_volume_num_voxels.restype = i64
    lib.ofx_set_graph.argtypes = [vp, vp, i32, f64, i32, vp]
    lib.ofx_set_transforms.argtypes = [vp, vp, vp, vp]
    lib.ofx_prepare_frame.argtypes = [vp, vp, vp, i32, i32, C.POINTER(C.c_float), vp]
    lib.ofx_stage_frame.argtypes = [vp, vp, vp, vp, vp, i32, i32, C.POINTER(C.c_float), vp]
    lib.ofx_integrate.argtypes = [vp, vp, vp, i32, i32, C.POINTER(C.c_float), f64, i32, vp, vp]
    lib.ofx_integrate_warped.argtypes = [vp, vp, vp, vp, vp, i32, i32, C.POINTER(C.c_float), f64, vp, vp]
    lib.ofx_volume_world_points.argtypes = [vp, vp, vp]
    lib.ofx_volume_skin.argtypes = [vp, vp, vp, vp, vp]
    lib.ofx_volume_deform.argtypes = [vp, vp, vp, vp]
    lib.ofx_pack_frame.argtypes = [vp, i32, i32, i32, vp, vp, vp]
    lib.ofx_skin.argtypes = [vp, i64, vp, i32, f64, i32, vp, vp, vp, vp]
    lib.ofx_warp_points.argtypes = [vp, i64, vp, vp, vp, i32, vp, vp, vp, vp]
    lib.ofx_peer_copy.argtypes = [vp, vp, i64, i32, vp]
    lib.ofx_multicast_copy.argtypes = [vp, vp, i64, i32, vp]
    lib.ofx_check_visibility.argtypes = [vp, i64, vp, i32, i32, C.POINTER(C.c_float), f64, vp, vp, vp]
    lib.ofx_nearest_node.argtypes = [vp, i64, vp, i32, vp, vp, vp]
    lib.ofx_pixel_anchors_euclidean.argtypes = [vp, i32, vp, i32, i32, C.c_float, vp, vp, vp]
    lib.ofx_argmin_norm.argtypes = [vp, i64, vp, i32, vp, vp, vp]
    lib.ofx_greedy_insert.argtypes = [vp, i32, vp, i32, vp, i32, f64, vp, vp, vp]
    lib.ofx_sample_nodes.argtypes = [vp, i32, vp, C.c_float, i32, vp, vp, vp, vp]
    lib.ofx_erode_mesh.argtypes = [vp, i32, i32, i32, i32, vp, vp]
    lib.ofx_truncated_region.argtypes = [vp, i32, i32, i32, f64, vp, vp]
    lib.ofx_mc_mesh.argtypes = [vp, vp, vp, i32, i32, i32, C.c_float, i32, i32, vp, vp, vp, vp, vp, vp]
    lib.ofx_compute_edges_geodesic.argtypes = [vp, vp, i32, vp, i32, vp, i32, i32, C.c_float, vp, vp, vp, i32, i32, vp, vp]
    lib.ofx_compute_edges_geodesic_batch.argtypes = [vp, vp, i32, vp, i32, vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.ofx_edge_weights_host.argtypes = [vp, vp, i32, i32, C.c_float, vp]
    lib.ofx_node_and_edge_clean_up.argtypes = [vp, i32, i32, vp, vp]
    lib.ofx_compute_clusters.argtypes = [vp, i32, i32, vp, vp, vp, vp]
    lib.ofx_compute_clusters_noalloc.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp]
    lib.ofx_upsample_index.argtypes = [vp, i32, vp, i32, vp, vp]
    lib.ofx_compute_edges_geodesic_host.argtypes = [vp, vp, i32, vp, i32, vp, i32, i32, C.c_float, vp, vp, vp, vp, i32, i32, i32]
    lib.ofx_node_and_edge_clean_up_host.argtypes = [vp, i32, i32, vp]
    lib.ofx_compute_clusters_host.argtypes = [vp, i32, i32, vp, vp, vp]
    lib.ofx_downsample_nodes_host.argtypes = [vp, i32, f64, vp, vp, vp]
    for name in SYMBOLS:
        getattr(lib, name)      # AttributeError here = the .so is older than include/ofx.h
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise OfxError(load().ofx_last_error().decode("utf-8", "replace"))


def require_cuda():
    if not torch.cuda.is_available():
        raise OfxError("no CUDA device: occlusionfusion_b200 runs its hot path only on the GPU "
                       "(no CPU fallback by design)")


_RAW_STREAM = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device=None):
    """cudaStream_t of torch's current stream on `device` (the raw getter: torch.cuda.current_stream() costs ~8 us of Python
    per call, and the frame loop asks several times per frame)."""
    if _RAW_STREAM is not None:
        idx = device.index if device is not None and getattr(device, "index", None) is not None else torch.cuda.current_device()
        return C.c_void_p(_RAW_STREAM(idx))
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return C.c_void_p(0)
    assert t.is_cuda and t.is_contiguous(), "expected a contiguous CUDA tensor"
    return C.c_void_p(t.data_ptr())


def intr4(cam_intr):
    """{fx, fy, cx, cy} cast to float32 as cam2pix does (tsdf.py:357-359)."""
    k = np.asarray(cam_intr, dtype=np.float64)
    if k.shape == (3, 3):
        k = np.array([k[0, 0], k[1, 1], k[0, 2], k[1, 2]])
    k32 = k.astype(np.float32)
    return (C.c_float * 4)(*[float(x) for x in k32])


def to_device(a, dtype, device):
    """numpy / tensor -> contiguous CUDA tensor of `dtype` (no copy when already there)."""
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=dtype).contiguous()
    return torch.from_numpy(np.ascontiguousarray(a)).to(device=device, dtype=dtype).contiguous()
