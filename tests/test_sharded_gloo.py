"""CPU, world_size 2 over gloo: the multi-GPU host path (graph broadcast, per-frame packet
broadcast, counter all-gather, volume gather with the dealt-block plane mapping).  The shard
compute is an oracle-backed stand-in (tests may use the oracle), so what is checked here is the
N>1 plumbing: two shards driven through ShardedFusion must reproduce the whole-volume oracle."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleShard:
    """shard backend with the interface of occlusionfusion_b200.sharded.CudaShard, on numpy."""

    def __init__(self, O, w, planes):
        self.O, self.w, self.planes = O, w, planes
        Dx, Dy, Dz = w.dims
        full = O.vox2world(w.origin, w.voxel_size, w.dims).reshape(Dx, Dy, Dz, 3)
        self.pts = np.ascontiguousarray(full[planes]).reshape(-1, 3)
        shape = (len(planes), Dy, Dz)
        self.tsdf, self.weight, self.colour = np.ones(shape, np.float32), np.zeros(shape, np.float32), np.zeros(shape, np.float32)

    def set_graph(self, nodes, coverage, K):
        self.skin = self.O.skin(self.pts, nodes.numpy(), K, coverage)

    def set_transforms(self, R, T):
        self.R, self.T = R.numpy().copy(), T.numpy().copy()

    def pack_frame(self, im):
        c, d = self.O.pack_frame(im)
        return torch.from_numpy(c), torch.from_numpy(d)

    def integrate(self, depth, colour, identity_warp, obs_weight):
        a, wt, v = self.skin
        if identity_warp:
            pts, valid = self.pts, np.ones(len(self.pts), bool)
        else:
            pts, valid = self.O.warp(self.pts, a, wt, v, self.R, self.T), v
        n = self.O.integrate(self.tsdf, self.weight, self.colour, pts, valid, depth.numpy(), colour.numpy(),
                             self.w.intr, self.w.trunc, obs_weight)
        return torch.tensor([n, int(v.sum()), 0, 0], dtype=torch.int64)

    def get_volume(self):
        return self.tsdf, self.colour, self.weight


def _worker(rank, world, port, x_block, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle as O
    from occlusionfusion_b200.sharded import ShardedFusion, default_x_block, plane_map
    from occlusionfusion_b200.synthetic import Workload
    O.set_threads(1)
    w = Workload("tiny", dim=[16, 24, 32], voxel=0.04, n_nodes=60, H=60, W=80)
    xb = default_x_block(w.dims[0], world) if x_block is None else x_block
    planes = plane_map(w.dims[0], world, rank, xb)
    sf = ShardedFusion(OracleShard(O, w, planes), w.dims, rank, world, xb, torch.device("cpu"))
    sf.set_graph(w.nodes if rank == 0 else None, w.coverage, w.K)
    assert sf.M == w.M and sf.K == w.K
    totals = []
    for f in range(3):
        if rank == 0:
            if f == 0:
                R, T = np.tile(np.eye(3, dtype=np.float32), (w.M, 1, 1)), np.zeros((w.M, 3), np.float32)
            else:
                e = w.estimate(f)
                R = e["node_rotations"]
                T = O.origin_translations(R, e["node_translations"], w.nodes)
            sf.fill_packet(w.frame(f), R, T)
        stats = sf.step(w.H, w.W, identity_warp=(f == 0))
        assert stats.shape == (world, 4)
        totals.append(int(stats[:, 0].sum()))
    full = sf.gather_volume()
    if rank == 0:
        np.savez(out_path, tsdf=full[0], colour=full[1], weight=full[2], totals=np.array(totals))
    else:
        assert full is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("x_block", [None, 8])
def test_two_shards_reproduce_whole_volume(tmp_path, oracle, x_block):
    from occlusionfusion_b200.synthetic import Workload
    port = 29600 + (os.getpid() % 300) + (0 if x_block is None else 1)
    out = str(tmp_path / "vol.npz")
    mp.spawn(_worker, args=(2, port, x_block, out), nprocs=2, join=True)
    got = np.load(out)
    w = Workload("tiny", dim=[16, 24, 32], voxel=0.04, n_nodes=60, H=60, W=80)
    ov = oracle.OracleVolume(w.dims, w.origin, w.voxel_size, w.intr, w.trunc)
    ov.set_graph(w.nodes, w.coverage, w.K)
    totals = []
    for f in range(3):
        if f:
            e = w.estimate(f)
            ov.set_transforms(e["node_rotations"], oracle.origin_translations(e["node_rotations"], e["node_translations"], w.nodes))
        totals.append(ov.integrate(w.frame(f), f == 0))
    assert list(got["totals"]) == totals and totals[1] > 0
    assert np.array_equal(got["weight"], ov.weight)
    assert np.array_equal(got["colour"], ov.colour)
    assert np.array_equal(got["tsdf"], ov.tsdf)
