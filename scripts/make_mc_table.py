#!/usr/bin/env python
"""Generate fluid_simulation_b200/csrc/mc_table.h: the marching-cubes case table, built from first principles.

Conventions.  Corner i of a cube sits at (i & 1, (i >> 1) & 1, (i >> 2) & 1) = (x, y, z); bit i of the case index
is set when that corner is INSIDE (value > level: an obstacle cell).  Edge e = 4 * axis + k runs along `axis`;
k = u + 2 * v where (u, v) are the cube coordinates of the edge on the two other axes, in increasing axis order.

Construction (no published table is consulted).  On each of the six faces the crossed edges are joined by
segments: two crossed edges give one segment; four (the ambiguous face: inside corners on a diagonal) give two
segments that each cut ONE inside corner off -- a rule that depends only on the face's own four corners, so the
two cubes sharing a face always agree and the surface is closed by construction.  Every crossed edge lies in
two faces, so the segments form closed loops; each loop becomes a triangle fan, oriented so that its normal
points from inside to outside (towards lower values: skimage's gradient_direction='descent', the default that
GUI/utils.py:17 uses).  Result: at most 5 triangles per cube, 820 in all.
"""
import itertools
import os

CORNER = [((i & 1), (i >> 1) & 1, (i >> 2) & 1) for i in range(8)]


def corner_index(p):
    return p[0] | (p[1] << 1) | (p[2] << 2)


EDGE_CORNER = []
for axis in range(3):
    others = [a for a in range(3) if a != axis]
    for k in range(4):
        p = [0, 0, 0]
        p[others[0]] = k & 1
        p[others[1]] = k >> 1
        q = list(p)
        q[axis] = 1
        EDGE_CORNER.append((corner_index(p), corner_index(q)))
EDGE_OF = {frozenset(c): e for e, c in enumerate(EDGE_CORNER)}


def faces():
    """Each face as its 4 corners in cyclic order."""
    out = []
    for axis in range(3):
        a, b = [x for x in range(3) if x != axis]
        for side in (0, 1):
            cyc = []
            for (u, v) in ((0, 0), (1, 0), (1, 1), (0, 1)):
                p = [0, 0, 0]
                p[axis] = side
                p[a] = u
                p[b] = v
                cyc.append(corner_index(p))
            out.append(cyc)
    return out


FACES = faces()


def segments(case):
    inside = [(case >> i) & 1 for i in range(8)]
    segs = []
    for cyc in FACES:
        crossed = []          # (edge, position j): edge between cyc[j] and cyc[j+1]
        for j in range(4):
            c0, c1 = cyc[j], cyc[(j + 1) % 4]
            if inside[c0] != inside[c1]:
                crossed.append((EDGE_OF[frozenset((c0, c1))], j))
        if len(crossed) == 2:
            segs.append((crossed[0][0], crossed[1][0]))
        elif len(crossed) == 4:
            # ambiguous face: cut each inside corner off on its own (the two edges that meet at it)
            for j in range(4):
                if inside[cyc[j]]:
                    e_prev = EDGE_OF[frozenset((cyc[(j - 1) % 4], cyc[j]))]
                    e_next = EDGE_OF[frozenset((cyc[j], cyc[(j + 1) % 4]))]
                    segs.append((e_prev, e_next))
    return segs


def loops(case):
    segs = segments(case)
    adj = {}
    for a, b in segs:
        adj.setdefault(a, []).append(b)
        adj.setdefault(b, []).append(a)
    assert all(len(v) == 2 for v in adj.values()), (case, adj)
    seen, out = set(), []
    for start in sorted(adj):
        if start in seen:
            continue
        loop, prev, cur = [start], None, start
        seen.add(start)
        while True:
            a, b = adj[cur]
            assert a != b                   # two faces never share two edges
            n = a if (prev is None or a != prev) else b
            if n == start:
                break
            loop.append(n)
            seen.add(n)
            prev, cur = cur, n
        out.append(loop)
    return out


def midpoint(e):
    a, b = EDGE_CORNER[e]
    return tuple((CORNER[a][i] + CORNER[b][i]) / 2.0 for i in range(3))


def oriented(case, loop):
    """Orient the loop so that its (Newell) normal points from the inside ends of its edges to the outside ends."""
    pts = [midpoint(e) for e in loop]
    n = [0.0, 0.0, 0.0]
    for i in range(len(pts)):
        p, q = pts[i], pts[(i + 1) % len(pts)]
        n[0] += (p[1] - q[1]) * (p[2] + q[2])
        n[1] += (p[2] - q[2]) * (p[0] + q[0])
        n[2] += (p[0] - q[0]) * (p[1] + q[1])
    s = 0.0
    for e in loop:
        a, b = EDGE_CORNER[e]
        if not (case >> a) & 1:
            a, b = b, a                     # a inside, b outside
        s += sum(n[i] * (CORNER[b][i] - CORNER[a][i]) for i in range(3))
    assert abs(s) > 1e-9, (case, loop)
    return loop if s > 0 else loop[::-1]


def build():
    table, total = [], 0
    for case in range(256):
        tris = []
        for loop in loops(case):
            lp = oriented(case, loop)
            for i in range(1, len(lp) - 1):
                tris.append((lp[0], lp[i], lp[i + 1]))
        assert len(tris) <= 5, (case, tris)
        total += len(tris)
        table.append(tris)
    return table, total


def main():
    table, total = build()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "fluid_simulation_b200", "csrc", "mc_table.h")
    with open(path, "w") as f:
        f.write("// mc_table.h -- GENERATED by scripts/make_mc_table.py (conventions and construction are described there).\n"
                "// Corner i = (i & 1, (i >> 1) & 1, (i >> 2) & 1); case bit i = corner i inside (value > level);\n"
                "// edge e = 4 * axis + k.  MC_TRI[case] = up to 5 triangles as edge triples, -1 terminated; normals\n"
                "// point towards lower values.  MC_NTRI[case] = the number of triangles.\n"
                "#pragma once\n#ifndef MC_TABLE_QUAL\n#define MC_TABLE_QUAL static const\n#endif\n")
        f.write("MC_TABLE_QUAL signed char MC_EDGE_CORNER[12][2] = {\n    ")
        f.write(", ".join("{%d, %d}" % c for c in EDGE_CORNER))
        f.write("\n};\nMC_TABLE_QUAL unsigned char MC_NTRI[256] = {\n")
        for r in range(0, 256, 32):
            f.write("    " + ", ".join(str(len(t)) for t in table[r:r + 32]) + ",\n")
        f.write("};\nMC_TABLE_QUAL signed char MC_TRI[256][16] = {\n")
        for case, tris in enumerate(table):
            flat = list(itertools.chain.from_iterable(tris))
            flat += [-1] * (16 - len(flat))
            f.write("    {" + ", ".join("%2d" % v for v in flat) + "},   // %d\n" % case)
        f.write("};\n")
    print("wrote", os.path.relpath(path), "max triangles", max(len(t) for t in table), "total", total)


if __name__ == "__main__":
    main()
