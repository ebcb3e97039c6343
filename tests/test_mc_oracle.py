"""CPU: the marching-cubes case table (tools/gen_mc_tables.py -> csrc/ofx_mc_tables.h) and the oracle's restatement
(oracle/ofx_oracle_mc.c).  There is no reference output to pin against (scikit-image absent), so the checks are the properties
a marching-cubes mesh must have: the committed table is what the generator derives, closed surfaces come out as closed
oriented manifolds, every interior edge of any field is shared by exactly two oppositely directed triangles, vertices sit at
the linear zero crossing of their grid edge."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _edges(f):
    return np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])


def test_table_is_what_the_generator_derives(tmp_path):
    path = os.path.join(ROOT, "occlusionfusion_b200", "csrc", "ofx_mc_tables.h")
    committed = open(path).read()
    src = open(os.path.join(ROOT, "tools", "gen_mc_tables.py")).read().replace(
        'OUT = os.path.join(', f'OUT = {str(tmp_path / "t.h")!r}\n_UNUSED = os.path.join(', 1)
    gen = tmp_path / "gen.py"
    gen.write_text(src)
    subprocess.check_call([sys.executable, str(gen)], stdout=subprocess.DEVNULL)
    assert open(tmp_path / "t.h").read() == committed


def test_table_uses_exactly_the_crossed_edges_of_every_case():
    """What csrc/ofx_mc.cu's vertex ownership rests on: the triangles of a case touch every grid edge of the cube whose
    end points lie on different sides of the level, and no other edge."""
    import re
    s = open(os.path.join(ROOT, "occlusionfusion_b200", "csrc", "ofx_mc_tables.h")).read()
    tris = [int(x) for x in re.findall(r"-?\d+", re.search(r"OFX_MC_TRIS\[256\]\s*=\s*\{(.*?)\};", s, re.S).group(1))]
    rows = re.findall(r"\{([^{}]*)\}", re.search(r"OFX_MC_EDGES\[256\]\[\d+\]\s*=\s*\{(.*)\};", s, re.S).group(1))
    edges = [[int(x) for x in re.findall(r"-?\d+", r)] for r in rows]
    assert len(tris) == 256 and len(edges) == 256
    for c in range(256):
        crossed = set()
        for e in range(12):
            axis, r = e >> 2, e & 3
            off = [0, 0, 0]
            off[1 if axis == 0 else 0] = r & 1
            off[1 if axis == 2 else 2] = r >> 1
            k0 = off[0] + 2 * off[1] + 4 * off[2]
            if ((c >> k0) & 1) != ((c >> (k0 + (1 << axis))) & 1):
                crossed.add(e)
        assert set(edges[c][:3 * tris[c]]) == crossed, c


def test_closed_surface_is_a_closed_oriented_manifold(oracle):
    n = 36
    x, y, z = np.meshgrid(*(np.arange(n),) * 3, indexing="ij")
    c = np.array([17.3, 18.1, 16.7])
    # two spheres joined by a neck: genus 0, one component
    d1 = np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) - 9.4
    d2 = np.sqrt((x - c[0] - 9) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) - 6.2
    vol = np.minimum(d1, d2).astype(np.float32)
    v, f, keys = oracle.marching_cubes(vol)
    e = _edges(f)
    und, counts = np.unique(np.sort(e, axis=1), axis=0, return_counts=True)
    assert (counts == 2).all() and len(np.unique(e, axis=0)) == len(e)       # two triangles per edge, opposite directions
    assert len(v) - len(und) + len(f) == 2
    a, b, cc = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    signed_volume = (a * np.cross(b, cc)).sum() / 6
    assert signed_volume > 0                                                  # normals from inside (negative) to outside
    assert np.array_equal(keys, np.unique(keys))
    # every vertex lies at the linear zero crossing of its edge: interpolating the field there gives 0
    lower, axis = keys // 3, keys % 3
    stride = np.array([n * n, n, 1])[axis]
    v0, v1 = vol.reshape(-1)[lower], vol.reshape(-1)[lower + stride]
    assert ((v0 < 0) != (v1 < 0)).all()
    t = v[np.arange(len(v)), axis] - np.stack([lower // (n * n), (lower // n) % n, lower % n], 1)[np.arange(len(v)), axis]
    assert np.abs(v0 + t * (v1 - v0)).max() < 1e-5


def test_random_fields_have_no_non_manifold_edges(oracle):
    for seed, dims in ((0, (14, 13, 15)), (1, (6, 20, 9)), (2, (11, 11, 11))):
        rng = np.random.default_rng(seed)
        vol = rng.normal(size=dims).astype(np.float32)
        v, f, _ = oracle.marching_cubes(vol)
        e = _edges(f)
        assert len(np.unique(e, axis=0)) == len(e)                            # no directed edge twice: consistent orientation
        _, counts = np.unique(np.sort(e, axis=1), axis=0, return_counts=True)
        assert set(counts.tolist()) <= {1, 2}                                 # 1 only on the volume boundary
        mask = rng.random(dims) < 0.5
        vm, fm, _ = oracle.marching_cubes(vol, mask)
        assert 0 < len(fm) < len(f)
    v, f, _ = oracle.marching_cubes(np.ones((5, 5, 5), np.float32))
    assert len(v) == 0 and f.shape == (0, 3)
