"""CPU: the oracle's restatement of the "next" rows against golden vectors produced
by RUNNING THE REFERENCE (oracle/make_golden_graph.py): its Python EDGraph.update / find_unreachable_nodes /
numba compute_truncated_region, and its C++ erode_mesh / sample_nodes compiled from /root/reference.
Integer / index / mask outputs: bit-exact."""
import numpy as np
import pytest

from conftest import load_golden


@pytest.fixture(scope="module")
def gg():
    return load_golden("ref_graph")


def test_erode_mesh_exact(oracle, gg):
    V = len(gg["mesh_verts"])
    for it, mn in ((1, 3), (3, 3), (2, 4), (0, 3)):
        got = oracle.erode_mesh(V, gg["mesh_faces"], it, mn)
        assert np.array_equal(got, gg[f"erode_{it}_{mn}"]), (it, mn)
    assert gg["erode_3_3"].sum() < gg["erode_1_3"].sum() < V       # the fixture does erode something


def test_sample_nodes_exact(oracle, gg):
    for use_ne in (1, 0):
        pos, idx = oracle.sample_nodes(gg["mesh_verts"], gg["erode_1_3"], float(gg["coverage"]), bool(use_ne))
        assert np.array_equal(idx, gg[f"sample_idx_{use_ne}"])
        assert np.array_equal(pos, gg[f"sample_pos_{use_ne}"])


def test_sample_nodes_summation_order_pinned(oracle, gg):
    """Pairs whose squared distance is <= coverage^2 under one float32 summation order and > under the other:
    the reference's compiled Eigen squaredNorm decides, the oracle has to agree on every pair."""
    pairs, want = gg["adv_pairs"], gg["adv_n_nodes"]
    assert 0 < (want == 1).sum() < len(want)
    for (p, q), n in zip(pairs, want):
        pos, idx = oracle.sample_nodes(np.stack([p, q]), np.ones(2, bool), float(gg["coverage"]), True)
        assert len(idx) == n


def test_nearest_node_and_unreachable_order(oracle, gg):
    dist, idx = oracle.nearest_node(gg["mesh_verts"], gg["upd_nodes_before"])
    # golden distances come from scipy's float64 tree cast to float32 (pykdtree shim): 1 ulp
    assert np.abs(dist - gg["upd_nearest_dist"]).max() <= 1e-7
    nd = gg["upd_nodes_before"]
    assert np.array_equal(idx, np.argmin(((gg["mesh_verts"][:, None, :].astype(np.float64) - nd[None]) ** 2).sum(-1), 1))
    unreachable = np.where(dist > 2 * float(gg["coverage"]))[0]               # warpfield.py:469
    assert set(unreachable.tolist()) == set(gg["upd_cand"].tolist())


def test_greedy_update_exact(oracle, gg):
    for cand, after in ((gg["upd_cand"], gg["upd_nodes_after"]), (gg["upd2_cand"], gg["upd2_nodes_after"])):
        before = gg["upd_nodes_before"]
        acc = oracle.greedy_update(gg["mesh_verts"], cand, before, float(gg["coverage"]))
        assert len(acc) == len(after) - len(before) > 0
        assert np.array_equal(gg["mesh_verts"][acc], after[len(before):])
    assert len(oracle.greedy_update(gg["mesh_verts"], np.zeros(0, np.int32), gg["upd_nodes_before"], 0.05)) == 0


def test_truncated_region_exact(oracle, gg):
    k4 = load_golden("ref_k4")
    n_true = 0
    for key in ("tsdf_0", "tsdf_2"):
        for md in (1.2, 0.5):
            got = oracle.truncated_region(k4[key], md)
            want = np.unpackbits(gg[f"trunc_{key}_{md}"])[:got.size].astype(bool).reshape(got.shape)
            assert np.array_equal(got, want), (key, md)
            n_true += int(want.sum())
    t = gg["trunc_smooth_in"]
    for md in (1.2, 0.3):
        got = oracle.truncated_region(t, md)
        want = np.unpackbits(gg[f"trunc_smooth_{md}"])[:got.size].astype(bool).reshape(got.shape)
        assert np.array_equal(got, want), md
        n_true += int(want.sum())
    assert n_true > 0
    assert np.unpackbits(gg["trunc_smooth_0.3"]).sum() < np.unpackbits(gg["trunc_smooth_1.2"]).sum()
