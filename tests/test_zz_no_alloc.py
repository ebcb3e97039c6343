"""GPU: ofx_compute_clusters_noalloc (include/ofx.h) -- the allocation-free variant of ofx_compute_clusters whose label
scratch is a file-scope device array -- called through the C ABI, twice in a row on different inputs, against the
allocating entry point (exact: the outputs are integers).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_NODES = 65536          # OFX_NOALLOC_MAX_NODES, include/ofx.h


def _random_edges(rng, M, E, p_stop=0.25):
    """(M, E) int32 neighbour lists: random node ids, each row cut by a -1 at a random place (graph_proc.cpp:455)."""
    e = rng.integers(0, M, size=(M, E)).astype(np.int32)
    e[np.arange(M) // 7 % 3 == 0] = -1                     # whole rows without neighbours: clusters of one
    cut = rng.random((M, E)) < p_stop
    first = np.where(cut.any(1), cut.argmax(1), E)
    e[np.arange(E)[None, :] >= first[:, None]] = -1
    # a few far-apart components: node ids of the upper half only point into the upper half
    half = M // 2
    lo = e[:half]
    lo[lo >= half] -= half
    hi = e[half:]
    hi[(hi >= 0) & (hi < half)] += half
    return np.ascontiguousarray(e)


def _run(lib, torch, edges, noalloc):
    from occlusionfusion_b200 import _capi
    M, E = edges.shape
    dev = torch.device("cuda", 0)
    e = torch.from_numpy(edges).to(dev)
    clusters = torch.full((M,), -7, dtype=torch.int32, device=dev)
    sizes = torch.full((M,), -7, dtype=torch.int32, device=dev)
    n = torch.full((1,), -7, dtype=torch.int32, device=dev)
    status = torch.full((1,), -7, dtype=torch.int32, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    if noalloc:
        rc = lib.ofx_compute_clusters_noalloc(p(e), M, E, p(clusters), p(sizes), p(n), p(status), stream)
    else:
        rc = lib.ofx_compute_clusters(p(e), M, E, p(clusters), p(sizes), p(n), stream)
    _capi.check(rc)
    torch.cuda.synchronize(dev)
    k = int(n.item())
    return clusters.cpu().numpy(), sizes.cpu().numpy()[:k], k, int(status.item())


def test_noalloc_clusters_equal_allocating_entry_point_twice():
    import torch
    from occlusionfusion_b200 import _capi
    assert torch.cuda.is_available()
    lib = _capi.load()
    rng = np.random.default_rng(20261105)
    with torch.cuda.device(0):
        # two calls in a row: the second is smaller, so stale labels of the first lie beyond (and inside) its range
        for M, E in [(3001, 8), (517, 4)]:
            edges = _random_edges(rng, M, E)
            c_ref, s_ref, k_ref, _ = _run(lib, torch, edges, noalloc=False)
            c, s, k, status = _run(lib, torch, edges, noalloc=True)
            assert status == 0
            assert k == k_ref and k > 2
            assert np.array_equal(c, c_ref) and np.array_equal(s, s_ref)
            assert int(s.sum()) == M
        # bad edge: reported through *status, no exception, no host read inside the call
        bad = np.array([[1, -1], [0, -1], [9, -1]], np.int32)
        _, _, _, status = _run(lib, torch, bad, noalloc=True)
        assert status == 3                                   # node 2 (+1)
        # and the next call is clean again
        edges = _random_edges(rng, 64, 3)
        c_ref, s_ref, k_ref, _ = _run(lib, torch, edges, noalloc=False)
        c, s, k, status = _run(lib, torch, edges, noalloc=True)
        assert status == 0 and k == k_ref and np.array_equal(c, c_ref) and np.array_equal(s, s_ref)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    from occlusionfusion_b200 import _capi
    lib = _capi.load()
    with torch.cuda.device(0):
        dev = torch.device("cuda", 0)
        M = MAX_NODES + 1
        e = torch.full((M, 1), -1, dtype=torch.int32, device=dev)
        out = torch.zeros((M,), dtype=torch.int32, device=dev)
        small = torch.zeros((2,), dtype=torch.int32, device=dev)
        p = lambda t: C.c_void_p(t.data_ptr())
        rc = lib.ofx_compute_clusters_noalloc(p(e), M, 1, p(out), p(out), p(small), p(small[1:]),
                                              C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
        assert rc != 0
        msg = lib.ofx_last_error()
        msg = msg.decode() if isinstance(msg, bytes) else str(msg)
        assert "exceed the static scratch" in msg
        torch.cuda.synchronize(dev)
