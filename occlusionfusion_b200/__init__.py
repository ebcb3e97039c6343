"""occlusionfusion_b200 -- B200-native (sm_100a) implementation of OcclusionFusion's non-rigid
fusion hot path: per-voxel warp-field evaluation fused with projective TSDF / weight / colour
integration, behind the reference's own Python interface.

    from occlusionfusion_b200 import TSDFVolume, WarpField, EDGraph

The compute lives in lib/libofx.so (C ABI: include/ofx.h; build: python -m occlusionfusion_b200.build).
There is no CPU fallback.
"""
from .embedded_deformation_graph import EDGraph
from .tsdf import TSDFVolume
from .warpfield import WarpField

__all__ = ["TSDFVolume", "WarpField", "EDGraph"]
