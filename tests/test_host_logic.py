"""CPU: host-side logic of the drop-in classes against the reference's golden vectors."""
import types

import numpy as np
import pytest

from conftest import load_golden


def test_volume_geometry_matches_reference(golden):
    from occlusionfusion_b200 import TSDFVolume
    _, g = golden
    K3 = np.eye(3)
    K3[0, 0], K3[1, 1], K3[0, 2], K3[1, 2] = g["intr"]
    fopt = types.SimpleNamespace(voxel_dim=[int(d) for d in g["voxel_dim"]])
    bnds, dims, vs = TSDFVolume.volume_geometry(tuple(g["bbox"]), float(g["max_depth"]), K3, fopt)
    assert np.array_equal(dims, g["voxel_dim"])
    assert vs == float(g["voxel_size"])
    assert np.array_equal(bnds[:, 0].astype(np.float32), g["vol_origin"])


def test_voxel_size_option():
    from occlusionfusion_b200 import TSDFVolume
    K3 = np.array([[525.0, 0, 319.5], [0, 525.0, 239.5], [0, 0, 1]])
    bnds, dims, vs = TSDFVolume.volume_geometry((0, 0, 640, 480), 2.0, K3, types.SimpleNamespace(voxel_size=0.01))
    assert vs == 0.01 and list(dims) == [int(np.ceil((bnds[i, 1] - bnds[i, 0]) / 0.01)) for i in range(3)]


def test_origin_translations_bit_identical_to_reference_loop(golden):
    """update_transformations uses one batched matmul; it must return the bits of the reference's
    per-node Python loop (warpfield.py:407-408), which the golden T arrays were produced with."""
    from occlusionfusion_b200.warpfield import _rot_apply
    _, g = golden
    for f in range(1, len(g["frames"])):
        R, t, nodes = g[f"R_{f}"], g[f"t_{f}"], g["nodes"]
        T = -_rot_apply(R, nodes) + nodes + t
        assert T.dtype == g[f"T_{f}"].dtype and np.array_equal(T, g[f"T_{f}"])
    rng = np.random.default_rng(1)
    for dt in (np.float32, np.float64):
        R, p = rng.normal(size=(3000, 3, 3)).astype(dt), rng.normal(size=(3000, 3)).astype(dt)
        assert np.array_equal(_rot_apply(R, p), np.array([R[i] @ p[i] for i in range(len(p))]))


def test_edgraph_update_needs_the_device():
    """Node insertion (embedded_deformation_graph.py:516-542) runs in libofx.so; without a GPU only the trivial
    early-outs work and everything else fails loudly (no CPU fallback).  Parity: tests/test_z_graph_gpu.py."""
    import torch
    from occlusionfusion_b200 import EDGraph
    from occlusionfusion_b200._capi import OfxError
    nodes = np.array([[0, 0, 0], [1, 0, 0]], np.float32)
    g = EDGraph(nodes=nodes, graph_generation_parameters={"node_coverage": 0.5})
    verts = np.array([[0.1, 0, 0], [2, 0, 0]], np.float32)
    assert g.update(verts, None, []) is False                       # :513-514
    if not torch.cuda.is_available():
        with pytest.raises(OfxError):
            g.update(verts, None, [0, 1])
        with pytest.raises(OfxError):
            g.erode_mesh(verts, np.zeros((1, 3), np.int32))


def test_graph_bin_formats_match_reference_bytes(tmp_path):
    """utils/utils.py:205-321: the files our writers produce are byte-identical to the ones the reference's own
    save_graph_* wrote (golden, oracle/make_golden_graph.py) and load back to the same arrays."""
    from conftest import load_golden
    from occlusionfusion_b200 import formats as F
    g = load_golden("ref_graph")
    for kind in ("nodes", "edges", "edges_weights", "node_deformations", "clusters"):
        a = g[f"fmt_{kind}_in"]
        path = str(tmp_path / (kind + ".bin"))
        getattr(F, "save_graph_" + kind)(path, a)
        assert open(path, "rb").read() == g[f"fmt_{kind}_bytes"].tobytes(), kind
        b = getattr(F, "load_graph_" + kind)(path)
        assert b.dtype == a.dtype and np.array_equal(b, a), kind
    with pytest.raises(AssertionError):
        F.load_graph_nodes(str(tmp_path / "missing.bin"))
    with pytest.raises(AssertionError):
        F.save_graph_nodes(str(tmp_path / "bad.bin"), np.zeros((3, 2), np.float32))


def test_plane_map_partitions_volume():
    from occlusionfusion_b200.sharded import default_x_block, plane_map
    for Dx, G, xb in [(512, 8, 4), (512, 8, 64), (40, 2, 4), (40, 2, 20), (24, 4, 2)]:
        planes = [plane_map(Dx, G, r, xb) for r in range(G)]
        allp = np.sort(np.concatenate(planes))
        assert np.array_equal(allp, np.arange(Dx))
        assert all(len(p) == Dx // G for p in planes)
    assert default_x_block(512, 8) == 4 and default_x_block(20, 2) == 10


def test_frame_packet_layout():
    import torch
    from occlusionfusion_b200.sharded import FramePacket
    p = FramePacket(M=7, H=6, W=10)
    buf = p.alloc("cpu")
    v = p.views(buf)
    assert v["R"].shape == (7, 3, 3) and v["T"].shape == (7, 3) and v["colour"].shape == (6, 10)
    assert p.o_colour % 4 == 0 and p.nbytes == 4 * buf.numel()
    v["T"].fill_(2.0); v["depth"].fill_(3.0)
    assert float(buf.sum()) == 2.0 * 21 + 3.0 * 60


def test_synthetic_workload_is_deterministic():
    from occlusionfusion_b200.synthetic import Workload
    a, b = Workload("tiny"), Workload("tiny")
    assert np.array_equal(a.nodes, b.nodes) and np.array_equal(a.frame(2), b.frame(2))
    ea, eb = a.estimate(3), b.estimate(3)
    assert np.array_equal(ea["node_rotations"], eb["node_rotations"])
    R = ea["node_rotations"].astype(np.float64)
    assert np.abs(R @ R.transpose(0, 2, 1) - np.eye(3)).max() < 1e-6
    w = Workload("c1")
    assert w.dims == (256, 256, 256) and abs(w.voxel_size - 0.01) < 1e-12 and w.M == 1000
    w = Workload("c2-dense")
    assert w.dims == (512, 512, 512) and abs(w.voxel_size - 0.0025) < 1e-12
