"""Asset pipeline: OBJ -> hull tables -> simulation.

CPU (`-m "not gpu"`): the library's OBJ reader (host code in libmadrona_b200.so, csrc/obj_hull.hpp) and the
half-edge tables built from its output must equal, bit for bit, what the unmodified reference's
PhysicsLoader::loadHullFromDisk makes of the same file -- against the committed fixtures everywhere
(tests/golden/assets_ref.npz, tests/golden/make_asset_golden.py) and against the compiled reference where it is
available. The pile of non-cube hulls is stepped by the oracle port against the compiled reference and by the
device source compiled for the host (tests/emu) against the port. GPU: the same pile through the CUDA step."""
import ctypes
import os

import numpy as np
import pytest

from gpu_ecs_madrona_b200 import assets, phys, scenes
from tests.golden import make_asset_golden as mag
from tests.util import assert_state_equal, bits, load_golden

REF_SHAPES = [n for n in scenes.HULL_SHAPES if n not in mag.TRIANGULAR]


def obj_path(name):
    return os.path.join(mag.HULLS, name + ".obj")


def polys_of(half_edges, polys):
    """Polygon index lists of a half-edge mesh, in the order the source listed them: a face's entry half-edge
    is the last one made for it, so its walk starts at the polygon's last vertex."""
    out = []
    for entry in polys:
        walk, he = [], int(entry)
        while True:
            walk.append(int(half_edges[he][2]))
            he = int(half_edges[he][0])
            if he == int(entry):
                break
        out.append(walk[1:] + walk[:1])
    return out


def check_against(name, want, lib, emu_lib):
    verts, polys, lo, hi = assets.load_hull_obj(obj_path(name), lib)
    assert np.array_equal(bits(verts), bits(want["verts"])), "vertex order / values"
    assert np.array_equal(bits(np.array(lo + hi, np.float32)), bits(want["aabb"]))
    assert polys == polys_of(want["half_edges"], want["polys"])
    # the tables mdr_phys_create builds from the loader's output (same host code in the CUDA library)
    sc = scenes.collisions64(1, num_cubes=1)
    sc.objects[0] = assets.hull_object(obj_path(name), lib=lib)
    e = phys.PhysExecutor(sc, lib=emu_lib)
    got = e.get_hull(0)
    for k in ("half_edges", "edges", "polys"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(bits(got["planes"]), bits(want["planes"]))
    e.close()


@pytest.fixture(scope="module")
def host_lib():
    """The product library: its OBJ entry points are host code and run without a GPU."""
    return phys.load_library()


@pytest.mark.parametrize("name", REF_SHAPES)
def test_obj_loader_matches_reference_fixture(host_lib, emu_lib, name):
    g = load_golden("assets_ref")
    want = {k: g[f"{name}_{k}"] for k in ("verts", "half_edges", "edges", "polys", "planes", "aabb")}
    check_against(name, want, host_lib, emu_lib)


@pytest.mark.parametrize("name", REF_SHAPES)
def test_obj_loader_matches_compiled_reference(host_lib, emu_lib, name):
    if not os.path.exists(mag.REF_LIB):
        pytest.skip("compiled reference asset loader not available on this machine")
    check_against(name, mag.ref_load(obj_path(name)), host_lib, emu_lib)


def test_obj_loader_random_files_match_compiled_reference(host_lib, tmp_path):
    """Seeded variants: other vertex permutations, number formats, line ends, `pos/uv/normal` corners."""
    if not os.path.exists(mag.REF_LIB):
        pytest.skip("compiled reference asset loader not available on this machine")
    shapes = mag.shapes()
    styles = ["g", "repr", "exp", "negzero"]
    for seed in range(24):
        name = REF_SHAPES[seed % len(REF_SHAPES)]
        v, f, _ = shapes[name]
        text = mag.obj_text(name, v, f, 1000 + seed, style=styles[seed % 4], crlf=seed % 3 == 0,
                            extras=seed % 2 == 0)
        if seed % 5 == 0:
            # one shared normal / uv for every corner: the merge key grows, the result must not change
            lines = text.splitlines(keepends=True)
            nl = "\r\n" if lines[0].endswith("\r\n") else "\n"
            body = []
            for ln in lines:
                if ln.startswith("f "):
                    ln = "f " + " ".join(c + "/1/1" for c in ln.split()[1:]) + nl
                body.append(ln)
            body.insert(2, "vn 0 0 1" + nl)
            body.insert(2, "vt 0.5 0.25" + nl)
            text = "".join(body)
        path = tmp_path / f"{name}_{seed}.obj"
        path.write_bytes(text.encode())
        want = mag.ref_load(str(path))
        verts, polys, lo, hi = assets.load_hull_obj(str(path), host_lib)
        assert np.array_equal(bits(verts), bits(want["verts"])), (name, seed)
        assert polys == polys_of(want["half_edges"], want["polys"]), (name, seed)
        assert np.array_equal(bits(np.array(lo + hi, np.float32)), bits(want["aabb"]))
        # text and file entry points agree
        v2, p2, _, _ = assets.load_hull_obj(text.encode(), host_lib)
        assert np.array_equal(bits(v2), bits(verts)) and p2 == polys


def test_obj_loader_triangular_and_errors(host_lib, emu_lib):
    verts, polys, lo, hi = assets.load_hull_obj(obj_path("tetra"), host_lib)
    assert verts.shape == (4, 3) and [len(p) for p in polys] == [3, 3, 3, 3]
    verts, polys, _, _ = assets.load_hull_obj(obj_path("rock"), host_lib)
    assert len(verts) == 13 and all(len(p) == 3 for p in polys) and len(polys) == 22
    bad = {
        "index 0": "v 0 0 0\nv 1 0 0\nv 0 1 0\nf 0 1 2\n",
        "index out of range": "v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 4\n",
        "relative index": "v 0 0 0\nv 1 0 0\nv 0 1 0\nf -3 -2 -1\n",
        "malformed number": "v 0 zero 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n",
        "leading plus": "v +1 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n",
        "tab separator": "v 0\t0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n",
        "mixed normals": "v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2 3\n",
        "no faces": "v 0 0 0\nv 1 0 0\nv 0 1 0\n",
        "empty": "# nothing\n\n",
        "two meshes": "o a\nv 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\no b\nv 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n",
    }
    for what, text in bad.items():
        with pytest.raises(phys.MadronaError) as ei:
            assets.load_hull_obj(text, host_lib)
        assert ei.value.code == phys.MDR_EINVAL, what
    with pytest.raises(phys.MadronaError):
        assets.load_hull_obj("/nonexistent/hull.obj", host_lib)
    # capacity protocol of the C entry point
    text = open(obj_path("cube"), "rb").read()
    nv, nw = ctypes.c_int32(3), ctypes.c_int32(100)
    v = np.zeros((8, 3), np.float32)
    w = np.zeros(100, np.uint32)
    rc = host_lib.mdr_load_hull_obj(text, len(text), v.ctypes.data, ctypes.byref(nv), w.ctypes.data,
                                    ctypes.byref(nw), None, None, 0)
    assert rc == phys.MDR_EOVERFLOW and nv.value == 8 and nw.value == 30
    # a closed mesh is required by the hull builder: an open one is refused at creation
    sc = scenes.collisions64(1, num_cubes=1)
    sc.objects[0].polys = sc.objects[0].polys[:-1]
    with pytest.raises(phys.MadronaError):
        phys.PhysExecutor(sc, lib=emu_lib)


def test_port_matches_compiled_reference_on_hull_pile(port_oracle, ref_oracle):
    sc = scenes.hull_pile(10)
    r = ref_oracle.Oracle(sc, "reference")
    d = ref_oracle.Oracle(sc, "reference_defined")     # locals defined: every world comparable (oracle/Makefile)
    p = port_oracle.Oracle(sc, "port")
    for st in range(40):
        r.step()
        d.step()
        p.step()
        det = p.indeterminate() == 0
        assert_state_equal(p.state(), r.state(), det, f"hull pile step {st}")
        assert_state_equal(p.state(), d.state(), None, f"hull pile step {st} (defined build)")
    assert det.sum() >= 5


def test_emu_matches_oracle_on_hull_pile(port_oracle, emu_lib):
    sc = scenes.hull_pile(6)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    for st in range(40):
        o.step()
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"hull pile step {st}")
    for k, obj in enumerate(sc.objects[:-1]):
        got, want = e.get_hull(k), o.get_hull(k)
        for name in ("half_edges", "edges", "polys"):
            assert np.array_equal(got[name], want[name]), (k, name)
        assert np.array_equal(bits(got["planes"]), bits(want["planes"])), k


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [32, 64, 128])
def test_cuda_matches_oracle_on_hull_pile(cuda_lib, port_oracle, threads):
    sc = scenes.hull_pile(40)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, block_threads=threads)
    for st in range(60):
        o.step(1, 8)
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"hull pile T={threads} step {st}")
    for col in (phys.COL_SUBSTEP_PREV, phys.COL_PRESOLVE_POS, phys.COL_PRESOLVE_VEL):
        assert np.array_equal(e.read(col), o.read(col)), col


@pytest.mark.gpu
def test_cuda_hull_tables_match_reference_fixture(cuda_lib):
    g = load_golden("assets_ref")
    sc = scenes.hull_pile(2)
    e = phys.PhysExecutor(sc)
    for k, name in enumerate(scenes.HULL_SHAPES):
        if name in mag.TRIANGULAR:
            continue
        got = e.get_hull(k)
        for t in ("half_edges", "edges", "polys"):
            assert np.array_equal(got[t], g[f"{name}_{t}"]), (name, t)
        assert np.array_equal(bits(got["planes"]), bits(g[f"{name}_planes"])), name


def _capacity_scene(n_sides, num_worlds):
    """A pile of n-gon prisms (2n vertices, n + 2 faces) and rocks over a plane."""
    v, f = mag.prism(n_sides, r=0.55, h=0.35)
    text = mag.obj_text(f"prism{n_sides}", v, f, 7, extras=False)
    sc = scenes.hull_pile(num_worlds, num_bodies=10, seed=23, spread=1.0)
    big = assets.hull_object(text, inv_mass=0.8, mu=0.4)
    sc.objects = [big, sc.objects[7]] + [sc.objects[-1]]          # the prism, the rock, the plane
    sc.obj_ids[:, 1:] = (np.arange(10)[None] + np.arange(num_worlds)[:, None]) % 2
    sc.obj_ids[:, 0] = 2
    return sc


def _barrel_scene(num_worlds, n=16, rings=4):
    """A pile of 64-vertex / 50-face barrels (two slots of the warp SAT) and rocks over a plane."""
    v, f = mag.barrel(n, rings)
    text = mag.obj_text(f"barrel{n}x{rings}", v, f, 9, extras=False)
    sc = scenes.hull_pile(num_worlds, num_bodies=8, seed=29, spread=1.0)
    big = assets.hull_object(text, inv_mass=0.7, mu=0.45)
    sc.objects = [big, sc.objects[7]] + [sc.objects[-1]]
    sc.obj_ids[:, 1:] = (np.arange(8)[None] + np.arange(num_worlds)[:, None]) % 2
    sc.obj_ids[:, 0] = 2
    return sc


def test_hull_at_capacity_and_beyond(port_oracle, emu_lib):
    """One slot of the warp SAT holds 32 vertices / faces, the big-hull variant two (MDR_MAX_HULL_VERTS / FACES = 64):
    32- and 64-vertex hulls step bit for bit, a 66-vertex one is refused at creation (no silent truncation)."""
    sc = _capacity_scene(16, 3)
    assert len(sc.objects[0].verts) == 32 and len(sc.objects[0].polys) == 18
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    for st in range(25):
        o.step()
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"32-vertex hulls step {st}")
    sc = _barrel_scene(2)
    assert len(sc.objects[0].verts) == 64 and len(sc.objects[0].polys) == 50
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    for st in range(15):
        o.step()
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"64-vertex hulls step {st}")
    with pytest.raises(phys.MadronaError):
        phys.PhysExecutor(_capacity_scene(33, 1), lib=emu_lib)


def test_port_matches_compiled_reference_at_hull_capacity(port_oracle, ref_oracle):
    for name, sc in (("32", _capacity_scene(16, 6)), ("64", _barrel_scene(6))):
        r = ref_oracle.Oracle(sc, "reference")
        d = ref_oracle.Oracle(sc, "reference_defined")
        p = port_oracle.Oracle(sc, "port")
        for st in range(30):
            r.step()
            d.step()
            p.step()
            det = p.indeterminate() == 0
            assert_state_equal(p.state(), r.state(), det, f"{name}-vertex hulls step {st}")
            assert_state_equal(p.state(), d.state(), None, f"{name}-vertex hulls step {st} (defined build)")
        assert det.sum() >= 3, name


@pytest.mark.gpu
def test_cuda_hull_at_capacity_and_beyond(cuda_lib, port_oracle):
    sc = _capacity_scene(16, 24)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc)
    for st in range(40):
        o.step(1, 8)
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"32-vertex hulls step {st}")
    with pytest.raises(phys.MadronaError):
        phys.PhysExecutor(_capacity_scene(33, 1))


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [32, 64])
def test_cuda_64_vertex_hulls(cuda_lib, port_oracle, threads):
    """hulls beyond 32 vertices / faces. 64-vertex, 50-face barrels through the two-slot warp SAT
    (its own kernel variant), bit for bit against the port (itself pinned on the compiled reference for this scene in
    test_port_matches_compiled_reference_at_hull_capacity), incl. the probed contact manifolds of one world."""
    sc = _barrel_scene(20)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, block_threads=threads)
    for st in range(40):
        o.step(1, 8)
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"64-vertex hulls step {st}")
    assert e.stats()[1].sum() > 0
    e.close(); o.close()


def test_obj_loader_dialect_matches_compiled_reference(host_lib, tmp_path):
    """Things real OBJ files carry that the reader must take exactly as the reference's does: other statement
    kinds (g, usemtl, mtllib, vp, l), a fourth coordinate / vertex colours on `v` lines, `vn` / `vt` lines
    nobody refers to, trailing blanks, blank lines, a file without a final newline, an `o` line first."""
    if not os.path.exists(mag.REF_LIB):
        pytest.skip("compiled reference asset loader not available on this machine")
    v, f, _ = mag.shapes()["frustum"]
    base = mag.obj_text("frustum", v, f, 77, extras=True).splitlines()
    variants = {}
    lines = []
    for ln in base:
        if ln.startswith("v "):
            ln = ln + " 1.0 0.25 0.5 0.75"             # w + colour
        if ln.startswith("f "):
            ln = ln + "   "                             # trailing blanks
        lines.append(ln)
        if ln.startswith("o "):
            lines += ["mtllib none.mtl", "g group0", "usemtl stone", "", "vp 0.1 0.2", "l 1 2"]
    lines.insert(5, "vn 0 0 1")
    lines.insert(6, "vt 0.5 0.5")
    variants["statements"] = "\n".join(lines) + "\n"
    variants["no_final_newline"] = "\n".join(base)
    variants["uv_only_corners"] = "\n".join(
        ("f " + " ".join(c + "/1" for c in ln.split()[1:])) if ln.startswith("f ") else ln for ln in base
    ).replace("s off", "vt 0.25 0.75\ns off") + "\n"
    variants["normal_only_corners"] = "\n".join(
        ("f " + " ".join(c + "//1" for c in ln.split()[1:])) if ln.startswith("f ") else ln for ln in base
    ).replace("s off", "vn 0 1 0\ns off") + "\n"
    for name, text in variants.items():
        path = tmp_path / (name + ".obj")
        path.write_bytes(text.encode())
        want = mag.ref_load(str(path))
        verts, polys, lo, hi = assets.load_hull_obj(str(path), host_lib)
        assert np.array_equal(bits(verts), bits(want["verts"])), name
        assert polys == polys_of(want["half_edges"], want["polys"]), name
        assert np.array_equal(bits(np.array(lo + hi, np.float32)), bits(want["aabb"])), name
