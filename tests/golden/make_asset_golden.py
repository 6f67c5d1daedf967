"""Generates the collision-hull fixtures of SURVEY §8f rank 4:

  gpu_ecs_madrona_b200/hulls/*.obj  convex test shapes as Wavefront OBJ (written here, committed; package data,
                               scenes.hull_pile loads them)
  tests/golden/assets_ref.npz  what the UNMODIFIED reference makes of each file:
                               PhysicsLoader::loadHullFromDisk (src/physics/physics_assets.cpp:205-252)
                               through oracle/_ref/libmadrona_assets_ref.so (oracle/Makefile, `assets_ref`)

Run in the build container (needs /root/reference and scipy): `python tests/golden/make_asset_golden.py`.
The OBJ texts deliberately use what the importer is sensitive to: `v` lines in an order other than the order
the faces mention them, an unused `v` line, duplicated `v` lines merged by value, `pos//normal`-free corners,
comments, `o` / `s` lines, CRLF line ends, exponents and signed zeros.
Fully triangular files (tetra, octa, rock) are NOT run through the reference: its loader dereferences a null
faceCounts array for them (see csrc/obj_hull.hpp); they are fixtures for the simulation parity tests only.
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))
HULLS = os.path.join(ROOT, "gpu_ecs_madrona_b200", "hulls")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libmadrona_assets_ref.so")

TRIANGULAR = ("tetra", "octa", "rock")


def ring(n, r, z, phase=0.0):
    a = phase + 2 * np.pi * np.arange(n) / n
    return np.stack([r * np.cos(a), r * np.sin(a), np.full(n, z)], -1)


def prism(n, r=0.5, h=0.5, top=1.0, phase=0.0):
    v = np.concatenate([ring(n, r, -h, phase), ring(n, r * top, h, phase)])
    f = [list(range(n - 1, -1, -1)), list(range(n, 2 * n))]
    f += [[k, (k + 1) % n, n + (k + 1) % n, n + k] for k in range(n)]
    return v, f


def barrel(n, rings, r=0.55, h=0.45):
    """`rings` rings of n vertices on an ellipsoidal profile (strictly convex): n * rings vertices,
    n * (rings - 1) quads + two n-gon caps."""
    zs = np.linspace(-h, h, rings)
    rad = r * np.sqrt(1.0 - 0.6 * (zs / h) ** 2)
    v = np.concatenate([ring(n, rad[k], zs[k]) for k in range(rings)])
    f = [list(range(n - 1, -1, -1)), list(range(n * (rings - 1), n * rings))]
    for k in range(rings - 1):
        f += [[k * n + i, k * n + (i + 1) % n, (k + 1) * n + (i + 1) % n, (k + 1) * n + i] for i in range(n)]
    return v, f


def pyramid(n, r=0.6, h=0.5, phase=0.0):
    v = np.concatenate([ring(n, r, -h * 0.5, phase), [[0.0, 0.0, h]]])
    f = [list(range(n - 1, -1, -1))] + [[k, (k + 1) % n, n] for k in range(n)]
    return v, f


def bipyramid(n, r=0.5, h=0.6):
    v = np.concatenate([ring(n, r, 0.0), [[0.0, 0.0, h]], [[0.0, 0.0, -h]]])
    f = [[k, (k + 1) % n, n] for k in range(n)] + [[(k + 1) % n, k, n + 1] for k in range(n)]
    return v, f


def convex_from_points(pts):
    """Faces of the convex hull of `pts` as CCW polygons (coplanar triangles merged)."""
    from scipy.spatial import ConvexHull
    hull = ConvexHull(pts)
    groups = {}
    for simplex, eq in zip(hull.simplices, hull.equations):
        key = tuple(np.round(eq, 6))
        groups.setdefault(key, set()).update(int(i) for i in simplex)
    centre = pts.mean(0)
    faces = []
    for key, idx in sorted(groups.items()):
        n = np.array(key[:3])
        idx = sorted(idx)
        c = pts[idx].mean(0)
        assert np.dot(n, c - centre) > 0
        u = pts[idx[0]] - c
        u /= np.linalg.norm(u)
        w = np.cross(n, u)
        ang = [np.arctan2(np.dot(pts[i] - c, w), np.dot(pts[i] - c, u)) for i in idx]
        faces.append([i for _, i in sorted(zip(ang, idx))])
    used = sorted({i for f in faces for i in f})
    assert used == list(range(len(pts))), "a point is inside the hull"
    return pts, faces


def dodecahedron(r=0.6):
    phi = (1 + 5 ** 0.5) / 2
    p = [(x, y, z) for x in (-1, 1) for y in (-1, 1) for z in (-1, 1)]
    for a in (-1 / phi, 1 / phi):
        for b in (-phi, phi):
            p += [(0, a, b), (a, b, 0), (b, 0, a)]
    p = np.array(p, np.float64)
    return convex_from_points(p * (r / np.linalg.norm(p[0])))


def rock(n=13, seed=5):
    rng = np.random.RandomState(seed)
    p = rng.normal(size=(n, 3))
    p /= np.linalg.norm(p, axis=1, keepdims=True)
    return convex_from_points(p * np.array([0.6, 0.5, 0.4]))


def fmt(x, style):
    x = float(np.float32(x))
    if style == "exp":
        return "%.8e" % x
    if x == 0.0 and style == "negzero":
        return "-0.0"
    return repr(x) if style == "repr" else "%.9g" % x


def obj_text(name, verts, faces, shuffle_seed, style="g", crlf=False, extras=True):
    """OBJ text with the `v` lines permuted, one unused `v`, one duplicated `v` (merged by the importer)."""
    rng = np.random.RandomState(shuffle_seed)
    verts = np.asarray(verts, np.float32)
    perm = rng.permutation(len(verts))            # line k holds vertex perm[k]
    where = np.argsort(perm)                      # vertex i is on line where[i] (0-based)
    lines = ["# %s: convex test hull, %d vertices, %d faces" % (name, len(verts), len(faces)), "o " + name]
    for k in perm:
        lines.append("v " + " ".join(fmt(c, style) for c in verts[k]))
    dup_of, dup_line = None, None
    if extras:
        lines.append("v 9.5 9.5 9.5")             # never referenced
        dup_of = int(rng.randint(len(verts)))
        lines.append("v  " + "  ".join(fmt(c, style) for c in verts[dup_of]))   # same value, double blanks
        dup_line = len(verts) + 2                 # 1-based index of the duplicate
    lines.append("s off")
    for fi, f in enumerate(faces):
        idx = [int(where[i]) + 1 for i in f]
        if extras and fi % 2 == 1:
            idx = [dup_line if i == dup_of else j for i, j in zip(f, idx)]
        lines.append("f " + " ".join(str(i) for i in idx))
    return ("\r\n" if crlf else "\n").join(lines) + ("\r\n" if crlf else "\n")


def shapes():
    out = {}
    out["cube"] = prism(4, r=0.5 * 2 ** 0.5, h=0.5, phase=np.pi / 4) + (dict(style="negzero"),)
    out["tri_prism"] = prism(3, r=0.6, h=0.45) + (dict(style="repr"),)
    out["hex_prism"] = prism(6, r=0.55, h=0.3) + (dict(style="exp"),)
    out["frustum"] = prism(4, r=0.7, h=0.4, top=0.5, phase=np.pi / 4) + (dict(crlf=True),)
    out["oct_frustum"] = prism(8, r=0.6, h=0.35, top=0.7) + (dict(),)
    out["pyramid"] = pyramid(4, phase=np.pi / 4) + (dict(),)
    out["dodeca"] = dodecahedron() + (dict(),)
    out["tetra"] = pyramid(3, r=0.7, h=0.7) + (dict(),)
    out["octa"] = bipyramid(4) + (dict(style="repr"),)
    out["rock"] = rock() + (dict(),)
    return out


def ref_load(path, cap=512):
    lib = ctypes.CDLL(REF_LIB)
    verts = np.zeros((cap, 3), np.float32); he = np.zeros((cap, 4), np.uint32)
    ed = np.zeros(cap, np.uint32); po = np.zeros(cap, np.uint32)
    pl = np.zeros((cap, 4), np.float32); ab = np.zeros(6, np.float32); cnt = np.zeros(4, np.int32)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
    lib.orc_load_hull_obj.argtypes = [ctypes.c_char_p, ctypes.c_int32] + [ctypes.c_void_p] * 7
    rc = lib.orc_load_hull_obj(path.encode(), cap, P(verts), P(he), P(ed), P(po), P(pl), P(ab), P(cnt))
    assert rc == 0
    return dict(verts=verts[:cnt[3]].copy(), half_edges=he[:cnt[0]].copy(), edges=ed[:cnt[1]].copy(),
                polys=po[:cnt[2]].copy(), planes=pl[:cnt[2]].copy(), aabb=ab)


def main():
    import subprocess
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "assets_ref"], check=True)
    os.makedirs(HULLS, exist_ok=True)
    out = {}
    for seed, (name, (v, f, kw)) in enumerate(sorted(shapes().items())):
        path = os.path.join(HULLS, name + ".obj")
        with open(path, "w", newline="") as fh:
            fh.write(obj_text(name, v, f, 100 + seed, **kw))
        if name in TRIANGULAR:
            continue
        for k, a in ref_load(path).items():
            out[f"{name}_{k}"] = a
    np.savez_compressed(os.path.join(HERE, "assets_ref.npz"), **out)
    print("wrote", len(out), "arrays for", sorted({k.split('_')[0] for k in out}))


if __name__ == "__main__":
    main()
