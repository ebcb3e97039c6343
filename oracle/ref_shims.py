"""TEST INFRASTRUCTURE ONLY -- import shims that let the *reference's own* Python modules
(`/root/reference/fusion_with_occlusion/tsdf.py`, `warpfield.py`) be imported in this
container, so that `oracle/make_golden.py` can run the reference CPU branch and write
golden vectors into `tests/golden/`.

Nothing in the product package may import this file.  It only works where
`/root/reference` exists (this build container); it never runs on the GPU box.

What is stubbed, and why (reference file:line of the import that needs it):
  open3d                      tsdf.py:9, warpfield.py:4        (visualisation only)
  skimage, skimage.measure    tsdf.py:16-17                    (marching cubes, not on the path)
  pykdtree.kdtree.KDTree      warpfield.py:11                  (pykdtree 1.3.4 is not installed:
                              replaced by scipy cKDTree, distances cast back to the query dtype
                              and indices to uint32 like pykdtree returns them)
  utils.utils / image_proc    warpfield.py:14                  (file helpers, not on the path)
  NeuralNRT._C                warpfield.py:15                  (compiled extension, not built)
  MVRegC                      NonRigidICP/model/geometry.py:3
  torch.cuda.init             tsdf.py:35 raises without a driver -> no-op
PyCUDA is absent, so tsdf.py:25-28 sets FUSION_GPU_MODE = 0 and the CPU branch
(tsdf.py:442-494, the parity contract) is selected by the reference itself.
"""
import os
import sys
import types

import numpy as np

REF_ROOT = os.environ.get("OFX_REFERENCE_ROOT", "/root/reference")


class _KDTreeShim:
    """pykdtree.kdtree.KDTree stand-in (call sites warpfield.py:51,103-104,463,532)."""

    def __init__(self, data, leafsize=16):
        from scipy.spatial import cKDTree
        self._data_dtype = np.asarray(data).dtype
        self._tree = cKDTree(np.asarray(data, dtype=np.float64), leafsize=leafsize)

    def query(self, points, k=1, **kw):
        dist, idx = self._tree.query(np.asarray(points, dtype=np.float64), k=k)
        return dist.astype(np.asarray(points).dtype), idx.astype(np.uint32)


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install():
    """Install the stubs and put the reference on sys.path. Idempotent."""
    if not os.path.isdir(REF_ROOT):
        raise RuntimeError(f"reference tree not found at {REF_ROOT}; golden vectors can only "
                           "be regenerated in the build container")
    if "open3d" not in sys.modules:
        _module("open3d")
    if "skimage" not in sys.modules:
        sk = _module("skimage")
        sk.measure = _module("skimage.measure")
        sk.io = _module("skimage.io")
    if "pykdtree" not in sys.modules:
        pk = _module("pykdtree")
        pk.kdtree = _module("pykdtree.kdtree", KDTree=_KDTreeShim)
    if "utils" not in sys.modules:
        u = _module("utils")
        u.utils = _module("utils.utils")
        u.image_proc = _module("utils.image_proc")
    if "NeuralNRT" not in sys.modules:
        n = _module("NeuralNRT")
        n._C = _module("NeuralNRT._C", compute_mesh_from_depth=None)
    if "MVRegC" not in sys.modules:
        _module("MVRegC")
    import torch
    torch.cuda.init = lambda: None
    for p in (os.path.join(REF_ROOT, "fusion_with_occlusion"), REF_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)


def import_reference():
    """Returns the reference modules (tsdf, warpfield)."""
    install()
    cwd = os.getcwd()
    try:
        os.chdir(os.path.join(REF_ROOT, "fusion_with_occlusion"))
        import tsdf as ref_tsdf            # noqa: E402
        import warpfield as ref_warpfield  # noqa: E402
    finally:
        os.chdir(cwd)
    return ref_tsdf, ref_warpfield
