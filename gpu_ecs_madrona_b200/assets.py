"""Collision assets: OBJ file -> hull object.

Host mirror of PhysicsLoader::loadHullFromDisk (src/physics/physics_assets.cpp:205-252):
``load_hull_obj`` calls the library's OBJ reader (``mdr_load_hull_obj`` / ``_file``, C++ in
csrc/obj_hull.hpp) and returns what an ``ObjectDef`` of type HULL takes. The half-edge tables are
built from that by ``mdr_phys_create`` (csrc/phys_assets.hpp), as HalfEdgeMesh::construct does.
Rigid-body metadata (inverse mass / inertia, friction) is the caller's, as in the reference, where
``loadObjects`` takes it next to the loaded mesh (physics_assets.cpp:254-396); ``solid_inertia``
offers the usual solid-polyhedron value for test scenes.
"""
import ctypes as C
import os

import numpy as np

from . import phys
from .scenes import HULL, ObjectDef


def load_hull_obj(source, lib=None):
    """`source`: path of an .obj file, or OBJ text (str / bytes containing a newline).
    Returns (verts [n, 3] f32, polys list of index lists, aabb_min, aabb_max)."""
    lib = lib or phys.load_library()
    nv, nw = C.c_int32(0), C.c_int32(0)
    aabb = np.zeros(6, np.float32)
    err = C.create_string_buffer(256)
    is_text = isinstance(source, bytes) or "\n" in source
    if is_text:
        text = source if isinstance(source, bytes) else source.encode()

        def call(v, w):
            return lib.mdr_load_hull_obj(text, len(text), v, C.byref(nv), w, C.byref(nw),
                                         aabb.ctypes.data, err, len(err))
    else:
        path = os.fspath(source).encode()

        def call(v, w):
            return lib.mdr_load_hull_obj_file(path, v, C.byref(nv), w, C.byref(nw),
                                              aabb.ctypes.data, err, len(err))
    rc = call(None, None)
    if rc != phys.MDR_OK:
        raise phys.MadronaError(rc, "load_hull_obj: " + err.value.decode())
    verts = np.zeros((nv.value, 3), np.float32)
    words = np.zeros(nw.value, np.uint32)
    rc = call(verts.ctypes.data, words.ctypes.data)
    if rc != phys.MDR_OK:
        raise phys.MadronaError(rc, "load_hull_obj: " + err.value.decode())
    polys, k = [], 0
    while k < len(words):
        n = int(words[k])
        polys.append([int(i) for i in words[k + 1:k + 1 + n]])
        k += n + 1
    return verts, polys, tuple(float(x) for x in aabb[:3]), tuple(float(x) for x in aabb[3:])


def solid_inertia(verts, polys, mass):
    """Diagonal of the inertia tensor of the uniform solid about the origin of object space
    (signed tetrahedra over a fan of every face; the reference keeps a Diag3x3, so objects are
    expected to be modelled in their principal frame)."""
    v = np.asarray(verts, np.float64)
    vol, second = 0.0, np.zeros(3)
    for p in polys:
        for k in range(1, len(p) - 1):
            a, b, c = v[p[0]], v[p[k]], v[p[k + 1]]
            d = float(np.dot(a, np.cross(b, c)))
            vol += d / 6.0
            # integral of x_i^2 over the tetrahedron (0, a, b, c)
            second += d / 60.0 * (a * a + b * b + c * c + a * b + a * c + b * c)
    dens = mass / vol
    s = second * dens
    return (s[1] + s[2], s[0] + s[2], s[0] + s[1])


def hull_object(source, inv_mass=1.0, mu=0.5, lib=None):
    """ObjectDef of a dynamic convex hull loaded from an OBJ (path or text)."""
    verts, polys, lo, hi = load_hull_obj(source, lib)
    inertia = solid_inertia(verts, polys, 1.0 / inv_mass)
    return ObjectDef(HULL, tuple(1.0 / i for i in inertia), inv_mass, mu, mu, lo, hi, verts, polys)
