"""Writes scenes/glass-box.{obj,mtl,camera}: a small closed room with two glass (Ni 1.5) icospheres, the test scene for the
reference's refraction / Fresnel branch (MTPC/pathTracing.cpp:13-27,77-114), which no shipped scene exercises (every
shipped material has Ni 1).  The files follow the grammar the reference's parsers accept (MTPC/sceneManagement.cpp:17-262:
one blank between tokens, triangles only, `f i/i/i` with v, vt and vn in lockstep), so the UNMODIFIED reference binary
(oracle/_ref/mcpt_ref) renders it; tests/golden/make_golden.py --glass turns those renders into the fixture
tests/golden/glass-box_image.npz.  Deterministic; 196 triangles (even: SURVEY.md 3A Q1); everything inside the Morton cube.

    python tests/golden/make_glass_scene.py
"""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def icosphere(center, radius, level=1):
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t), (t, 0, -1), (t, 0, 1),
         (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6), (7, 1, 8),
         (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    v = [np.array(p, float) / np.linalg.norm(p) for p in v]
    for _ in range(level):
        nf, mid = [], {}

        def m(a, b):
            k = (min(a, b), max(a, b))
            if k not in mid:
                p = v[a] + v[b]
                v.append(p / np.linalg.norm(p))
                mid[k] = len(v) - 1
            return mid[k]
        for a, b, c in f:
            ab, bc, ca = m(a, b), m(b, c), m(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    tris = []
    for a, b, c in f:
        # smooth (unit) vertex normals; positions on the sphere
        tris.append(([center + radius * v[i] for i in (a, b, c)], [v[i] for i in (a, b, c)]))
    return tris


def quad(p0, p1, p2, p3, n):
    n = np.array(n, float)
    P = [np.array(p, float) for p in (p0, p1, p2, p3)]
    return [([P[0], P[1], P[2]], [n, n, n]), ([P[0], P[2], P[3]], [n, n, n])]


def main():
    groups = []   # (material, triangles)
    groups.append(("Floor", quad((-1, 0, -1), (-1, 0, 1), (1, 0, 1), (1, 0, -1), (0, 1, 0))))
    groups.append(("Ceiling", quad((-1, 2, -1), (1, 2, -1), (1, 2, 1), (-1, 2, 1), (0, -1, 0))))
    groups.append(("BackWall", quad((-1, 0, -1), (1, 0, -1), (1, 2, -1), (-1, 2, -1), (0, 0, 1))))
    groups.append(("LeftWall", quad((-1, 0, -1), (-1, 2, -1), (-1, 2, 1), (-1, 0, 1), (1, 0, 0))))
    groups.append(("RightWall", quad((1, 0, -1), (1, 0, 1), (1, 2, 1), (1, 2, -1), (-1, 0, 0))))
    groups.append(("Light", quad((-0.3, 1.98, -0.3), (0.3, 1.98, -0.3), (0.3, 1.98, 0.3), (-0.3, 1.98, 0.3), (0, -1, 0))))
    groups.append(("Glass", icosphere(np.array([-0.42, 0.55, 0.15]), 0.42)))
    groups.append(("Glass", icosphere(np.array([0.45, 1.05, -0.25]), 0.38)))
    n_tris = sum(len(t) for _, t in groups)
    assert n_tris % 2 == 0
    out = os.path.join(ROOT, "scenes")
    with open(os.path.join(out, "glass-box.obj"), "w", newline="\n") as fo:
        vi = 0
        body = []
        for mat, tris in groups:
            body.append("usemtl %s" % mat)
            for pos, nrm in tris:
                for k in range(3):
                    fo.write("v %.6f %.6f %.6f\n" % tuple(pos[k]))
                for k in range(3):
                    fo.write("vn %.6f %.6f %.6f\n" % tuple(nrm[k]))
                for k in range(3):
                    fo.write("vt %.6f %.6f\n" % (0.0, 0.0))
                body.append("f %d/%d/%d %d/%d/%d %d/%d/%d" % (vi + 1, vi + 1, vi + 1, vi + 2, vi + 2, vi + 2, vi + 3, vi + 3, vi + 3))
                vi += 3
        fo.write("\n".join(body) + "\n")
    with open(os.path.join(out, "glass-box.mtl"), "w", newline="\n") as fo:
        def mtl(name, kd, ks=(0, 0, 0), ns=1, ni=1):
            fo.write("newmtl %s\nKd %g %g %g\nKs %g %g %g\nNs %g\nNi %g\n" % ((name,) + tuple(kd) + tuple(ks) + (ns, ni)))
        mtl("Floor", (0.75, 0.75, 0.75)); mtl("Ceiling", (0.75, 0.75, 0.75)); mtl("BackWall", (0.75, 0.75, 0.75))
        mtl("LeftWall", (0.63, 0.065, 0.05)); mtl("RightWall", (0.14, 0.091, 0.45))
        mtl("Light", (0, 0, 0))
        mtl("Glass", (0.05, 0.05, 0.05), (0.9, 0.9, 0.9), 200, 1.5)
    with open(os.path.join(out, "glass-box.camera"), "w", newline="\n") as fo:
        fo.write("eye 0 1 4.4\nlookat 0 1 3.4\nup 0 1 0\nfovy 32\nwidth 160\nheight 160\nmtlname Light 30 30 30\n")
    print("glass-box: %d triangles" % n_tris)


if __name__ == "__main__":
    main()
