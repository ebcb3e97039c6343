"""Generates occlusionfusion_b200/csrc/ofx_mc_tables.h: the marching-cubes case table used by the CUDA kernels
(csrc/ofx_mc.cu) and by the test-only CPU checker (which includes this header).  Usage: python tools/gen_mc_tables.py

The table is DERIVED, not transcribed: for each of the 256 sign patterns of a cube's corners the isosurface polygons
are built from the six faces.  On a face (seen from outside the cube, corners counter-clockwise) every maximal run of
inside corners is cut off by one directed chord from the edge where the boundary leaves the run to the edge where it
enters it; the rule depends on the face's four signs only, so the two cubes sharing a face draw the same chords and the
mesh is watertight, and a face with two separated inside corners (the ambiguous pattern) always separates them.  Every
crossed cube edge is the tail of one chord and the head of another (its two faces run along it in opposite directions),
so the chords close into oriented loops; a loop is triangulated as a fan.  Triangles are oriented with their normal
pointing from inside (value < level) to outside.

Conventions: corner c = x + 2*y + 4*z (bits of the corner's offset); edge e = 4*axis + (b0 + 2*b1) where (b0, b1) are
the corner offsets along the other two axes in increasing axis order; case index bit c = corner c is inside.
"""
import itertools
import os

import numpy as np

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "occlusionfusion_b200", "csrc",
                   "ofx_mc_tables.h")


def corner_id(p):
    return p[0] + 2 * p[1] + 4 * p[2]


def edge_id(p, q):
    """Edge between corners p and q (offset triples differing along one axis)."""
    axis = [i for i in range(3) if p[i] != q[i]]
    assert len(axis) == 1
    a = axis[0]
    others = [i for i in range(3) if i != a]
    return 4 * a + p[others[0]] + 2 * p[others[1]]


def edge_corners(e):
    a, r = divmod(e, 4)
    others = [i for i in range(3) if i != a]
    p = [0, 0, 0]
    p[others[0]], p[others[1]] = r & 1, r >> 1
    q = list(p)
    q[a] = 1
    return tuple(p), tuple(q)


def face_cycles():
    """The 6 faces as corner cycles, counter-clockwise seen from outside the cube."""
    faces = []
    for a in range(3):
        u, v = (a + 1) % 3, (a + 2) % 3                     # (u, v, a) is right-handed
        for s in (0, 1):
            cyc = []
            for cu, cv in ((0, 0), (1, 0), (1, 1), (0, 1)):
                p = [0, 0, 0]
                p[a], p[u], p[v] = s, cu, cv
                cyc.append(tuple(p))
            faces.append(cyc if s == 1 else cyc[::-1])
    return faces


def build_case(case, faces):
    inside = lambda p: (case >> corner_id(p)) & 1        # noqa: E731
    nxt = {}                                               # chord: tail edge -> head edge
    for cyc in faces:
        ins = [inside(p) for p in cyc]
        for k in range(4):
            # boundary leaves an inside run between corner k (inside) and k+1 (outside): chord tail
            if ins[k] and not ins[(k + 1) % 4]:
                tail = edge_id(cyc[k], cyc[(k + 1) % 4])
                # walk backwards to the start of this inside run: the edge where the boundary entered it
                j = k
                while ins[(j - 1) % 4] and (j - 1) % 4 != k:
                    j = (j - 1) % 4
                head = edge_id(cyc[(j - 1) % 4], cyc[j])
                assert tail not in nxt
                nxt[tail] = head
    loops, seen = [], set()
    for start in sorted(nxt):
        if start in seen:
            continue
        loop, e = [], start
        while e not in seen:
            seen.add(e)
            loop.append(e)
            e = nxt[e]
        assert e == start
        loops.append(loop)
    # triangulate every loop; a diagonal whose two crossings lie on one cube face would coincide with a diagonal of the
    # neighbouring cube (an edge with four triangles), so pick a triangulation without such diagonals
    edge_faces = {}
    for fi, cyc in enumerate(faces):
        for k in range(4):
            edge_faces.setdefault(edge_id(cyc[k], cyc[(k + 1) % 4]), set()).add(fi)

    def in_face(a, b):
        return bool(edge_faces[a] & edge_faces[b])

    def triangulations(poly):
        if len(poly) == 3:
            yield [tuple(poly)]
            return
        a, b = poly[0], poly[1]                              # the triangle on edge (poly[0], poly[1]) picks its third corner
        for k in range(2, len(poly)):
            left = [[]] if k == 2 else triangulations(poly[1:k + 1])
            for lt in left:
                right = [[]] if k == len(poly) - 1 else triangulations([poly[0]] + poly[k:])
                for rt in right:
                    yield [(a, b, poly[k])] + lt + rt

    tris = []
    for loop in loops:
        best, best_bad = None, None
        for cand in triangulations(loop):
            sides = {frozenset(p) for p in zip(loop, loop[1:] + loop[:1])}
            diag = {frozenset((t[i], t[(i + 1) % 3])) for t in cand for i in range(3)} - sides
            bad = sum(in_face(*tuple(d)) for d in diag)
            if best is None or bad < best_bad:
                best, best_bad = cand, bad
            if bad == 0:
                break
        tris += best
    return tris


def main():
    faces = face_cycles()
    table = [build_case(c, faces) for c in range(256)]
    # orientation check on the single-corner cases; flip everything if the normals point inwards
    mid = lambda e: np.mean(np.array(edge_corners(e), float), axis=0)      # noqa: E731
    flips = []
    for c in range(8):
        (t,) = table[1 << c]
        a, b, d = (mid(e) for e in t)
        corner = np.array([(c >> i) & 1 for i in range(3)], float)
        flips.append(np.dot(np.cross(b - a, d - a), np.full(3, 0.5) - corner) < 0)
    assert all(flips) or not any(flips)
    if flips[0]:
        table = [[(t[0], t[2], t[1]) for t in tris] for tris in table]
    # checks: crossings used exactly match sign changes; complementary cases mirror each other's edge sets
    for c in range(256):
        crossed = {e for e in range(12) if ((c >> corner_id(edge_corners(e)[0])) & 1) != ((c >> corner_id(edge_corners(e)[1])) & 1)}
        used = {e for t in table[c] for e in t}
        assert used == crossed, c
        # every undirected polygon edge inside the cube is shared by two triangles or lies on a face
    max_tris = max(len(t) for t in table)
    assert max_tris <= 5, max_tris
    lines = ["// GENERATED by tools/gen_mc_tables.py -- do not edit.  Marching-cubes case table shared by csrc/ofx_mc.cu and the",
             "// test-only CPU checker.  corner c = x + 2y + 4z; edge e = 4*axis + (b0 + 2*b1); case bit c = corner c inside",
             "// (value < level).  OFX_MC_TRIS[case] = number of triangles, OFX_MC_EDGES[case][3*t + k] = edge of corner k of",
             "// triangle t (normal from inside to outside).",
             "#ifndef OFX_MC_TABLES_H", "#define OFX_MC_TABLES_H", f"#define OFX_MC_MAX_TRIS {max_tris}",
             "#ifdef __CUDACC__", "#define OFX_MC_CONST static __device__ __constant__", "#else",
             "#define OFX_MC_CONST static const", "#endif",
             "OFX_MC_CONST unsigned char OFX_MC_TRIS[256] = {"]
    counts = [len(t) for t in table]
    for i in range(0, 256, 32):
        lines.append("    " + ", ".join(str(x) for x in counts[i:i + 32]) + ",")
    lines += ["};", f"OFX_MC_CONST signed char OFX_MC_EDGES[256][{3 * max_tris}] = {{"]
    for c in range(256):
        flat = [e for t in table[c] for e in t]
        flat += [-1] * (3 * max_tris - len(flat))
        lines.append("    {" + ", ".join(f"{x:2d}" for x in flat) + "},")
    lines += ["};", "#endif", ""]
    with open(OUT, "w") as f:
        f.write("\n".join(lines))
    print(f"wrote {OUT}: max {max_tris} triangles per cube, {sum(counts)} triangles over the 256 cases")


if __name__ == "__main__":
    main()
