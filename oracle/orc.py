"""TEST INFRASTRUCTURE ONLY: ctypes loader for the CPU checkers behind
oracle/oracle_api.h (the compiled reference `_ref/libmadrona_ref.so`, kind
"reference", and our restatement `libmadrona_port.so`, kind "port").
Imported only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libmadrona_ref.so")
# the same sources with automatic variables zero-initialised (oracle/Makefile): kind "reference_defined"
REF_DEFINED_SO = os.path.join(HERE, "_ref", "libmadrona_ref_defined.so")
PORT_SO = os.path.join(HERE, "libmadrona_port.so")

COL_BYTES = [8, 12, 16, 12, 24, 4, 4, 28, 28, 24, 12, 12, 4]
(COL_ENTITY, COL_POSITION, COL_ROTATION, COL_SCALE, COL_VELOCITY, COL_OBJECT_ID,
 COL_RESPONSE_TYPE, COL_SUBSTEP_PREV, COL_PRESOLVE_POS, COL_PRESOLVE_VEL,
 COL_EXT_FORCE, COL_EXT_TORQUE, COL_LEAF_ID) = range(13)


class OrcObject(C.Structure):
    _fields_ = [("type", C.c_int32), ("inv_inertia", C.c_float * 3),
                ("inv_mass", C.c_float), ("mu_s", C.c_float), ("mu_d", C.c_float),
                ("aabb_min", C.c_float * 3), ("aabb_max", C.c_float * 3),
                ("vert_offset", C.c_int32), ("num_verts", C.c_int32),
                ("poly_offset", C.c_int32), ("num_poly_words", C.c_int32)]


class OrcConfig(C.Structure):
    _fields_ = [("num_worlds", C.c_int32), ("bodies_per_world", C.c_int32),
                ("num_substeps", C.c_int32), ("max_contacts", C.c_int32),
                ("delta_t", C.c_float), ("gravity", C.c_float * 3)]


class OrcProbe(C.Structure):
    _fields_ = [("max_pairs", C.c_int32), ("num_pairs", C.c_int32),
                ("pairs", C.c_void_p), ("num_contacts", C.c_void_p),
                ("contacts", C.c_void_p), ("slot_boxes", C.c_void_p),
                ("leaf_aabbs", C.c_void_p), ("ref_rank", C.c_void_p)]


CONTACT_DTYPE = np.dtype([("ref_row", "<i4"), ("ref_pad", "<i4"), ("alt_row", "<i4"),
                          ("alt_pad", "<i4"), ("points", "<f4", (4, 4)),
                          ("num_points", "<i4"), ("normal", "<f4", (3,)),
                          ("lambda_n", "<f4", (4,))])
assert CONTACT_DTYPE.itemsize == 112


def build(kind):
    """(Re)build a checker with oracle/Makefile. `ref` needs /root/reference."""
    subprocess.run(["make", "-C", HERE, kind], check=True, capture_output=True)


def _path(kind):
    return {"reference": REF_SO, "reference_defined": REF_DEFINED_SO}.get(kind, PORT_SO)


def available(kind):
    return os.path.exists(_path(kind))


def pack_objects(objects):
    verts, words, arr = [], [], (OrcObject * len(objects))()
    voff = woff = 0
    for i, o in enumerate(objects):
        pw = o.poly_words()
        arr[i].type = o.type
        arr[i].inv_inertia = (C.c_float * 3)(*o.inv_inertia)
        arr[i].inv_mass, arr[i].mu_s, arr[i].mu_d = o.inv_mass, o.mu_s, o.mu_d
        arr[i].aabb_min = (C.c_float * 3)(*o.aabb_min)
        arr[i].aabb_max = (C.c_float * 3)(*o.aabb_max)
        arr[i].vert_offset, arr[i].num_verts = voff, len(o.verts)
        arr[i].poly_offset, arr[i].num_poly_words = woff, len(pw)
        verts.append(np.asarray(o.verts, np.float32).reshape(-1, 3))
        words.append(pw)
        voff += len(o.verts)
        woff += len(pw)
    v = np.ascontiguousarray(np.concatenate(verts) if verts else np.zeros((0, 3), np.float32))
    w = np.ascontiguousarray(np.concatenate(words) if words else np.zeros(0, np.uint32))
    if len(v) == 0:
        v = np.zeros((1, 3), np.float32)
    if len(w) == 0:
        w = np.zeros(1, np.uint32)
    return arr, v, w


class Oracle:
    """One checker instance over a `gpu_ecs_madrona_b200.scenes.Scene`."""

    def __init__(self, scene, kind="port"):
        path = _path(kind)
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} not built (make -C oracle)")
        self.lib = C.CDLL(path)
        L = self.lib
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_void_p] * 2 + [C.c_int32] + [C.c_void_p] * 10
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_step.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
        L.orc_read.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
        L.orc_write.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
        L.orc_probe_step.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
        L.orc_probe_step.restype = C.c_int32
        L.orc_flag_rebuild.argtypes = [C.c_void_p]
        L.orc_flag_rebuild_world.argtypes = [C.c_void_p, C.c_int32]
        L.orc_get_hull.argtypes = [C.c_void_p, C.c_int32] + [C.c_void_p] * 5
        L.orc_indeterminate.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_trace_rays.argtypes = [C.c_void_p, C.c_int32, C.c_int32] + [C.c_void_p] * 6
        L.orc_feeder.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32] + [C.c_void_p] * 4
        L.orc_set_joints.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
        L.orc_kind.restype = C.c_char_p
        self.kind = L.orc_kind().decode()
        self.scene = scene
        cfg = OrcConfig(scene.num_worlds, scene.bodies_per_world, scene.num_substeps,
                        scene.max_contacts, scene.delta_t, (C.c_float * 3)(*scene.gravity))
        objs, verts, words = pack_objects(scene.objects)
        c = np.ascontiguousarray
        self._keep = [c(scene.pos, np.float32), c(scene.rot, np.float32),
                      c(scene.scale, np.float32), c(scene.vel, np.float32),
                      c(scene.obj_ids, np.int32), c(scene.response, np.uint32),
                      c(scene.ext_force, np.float32), c(scene.ext_torque, np.float32)]
        self.h = L.orc_create(C.byref(cfg), objs, len(scene.objects),
                              verts.ctypes.data, words.ctypes.data,
                              *[a.ctypes.data for a in self._keep])
        if getattr(scene, "joints", None) is not None:
            j = c(scene.joints)
            assert j.dtype.itemsize == 84 and j.shape[0] == scene.num_worlds
            n = c(scene.joint_counts, np.int32)
            L.orc_set_joints(self.h, j.shape[1], n.ctypes.data, j.ctypes.data)

    def close(self):
        if self.h:
            self.lib.orc_destroy(self.h)
            self.h = None

    def step(self, num_steps=1, num_threads=1):
        self.lib.orc_step(self.h, num_steps, num_threads)

    def flag_rebuild(self, worlds=None):
        if worlds is None:
            self.lib.orc_flag_rebuild(self.h)
        else:
            for w in worlds:
                self.lib.orc_flag_rebuild_world(self.h, int(w))

    def read(self, col):
        W, N = self.scene.num_worlds, self.scene.bodies_per_world
        out = np.empty((W, N, COL_BYTES[col]), np.uint8)
        self.lib.orc_read(self.h, col, out.ctypes.data)
        return out

    def read_f32(self, col):
        return self.read(col).view(np.float32)

    def write(self, col, arr):
        arr = np.ascontiguousarray(arr)
        self.lib.orc_write(self.h, col, arr.ctypes.data)

    def state(self):
        """(pos[W,N,3], rot[W,N,4], vel[W,N,6]) after the last step."""
        return (self.read_f32(COL_POSITION), self.read_f32(COL_ROTATION),
                self.read_f32(COL_VELOCITY))

    def probe_step(self, world, max_pairs=4096):
        sc = self.scene
        N, S, MC = sc.bodies_per_world, sc.num_substeps, sc.max_contacts
        pairs = np.zeros((max_pairs, 2), np.int32)
        ncon = np.zeros(S, np.int32)
        contacts = np.zeros((S, MC), CONTACT_DTYPE)
        slot = np.zeros((N, 6), np.float32)
        leaf = np.zeros((N, 6), np.float32)
        rank = np.zeros(N, np.int32)
        p = OrcProbe(max_pairs, 0, pairs.ctypes.data, ncon.ctypes.data, contacts.ctypes.data,
                     slot.ctypes.data, leaf.ctypes.data, rank.ctypes.data)
        rc = self.lib.orc_probe_step(self.h, world, C.byref(p))
        if rc != 0:
            raise RuntimeError("probe overflow")
        return dict(pairs=pairs[:p.num_pairs].copy(), num_contacts=ncon,
                    contacts=[contacts[s, :ncon[s]].copy() for s in range(S)],
                    slot_boxes=slot, leaf_aabbs=leaf, ref_rank=rank)

    def trace_rays(self, world, origins, dirs, t_max):
        """BVH::traceRay for rays [n,3] in one world: (t[n], normal[n,3], row[n])."""
        o = np.ascontiguousarray(origins, np.float32)
        d = np.ascontiguousarray(dirs, np.float32)
        n = o.shape[0]
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(t_max, np.float32), (n,)))
        t = np.zeros(n, np.float32)
        nrm = np.zeros((n, 3), np.float32)
        row = np.zeros(n, np.int32)
        self.lib.orc_trace_rays(self.h, world, n, o.ctypes.data, d.ctypes.data, tm.ctypes.data,
                                t.ctypes.data, nrm.ctypes.data, row.ctypes.data)
        return t, nrm, row

    def feeder(self, world_offset, view_rows, camera_offset, y_scale):
        """(transforms[W,N,12], views[W,V,8]) of the renderer feeder's arithmetic."""
        W, N = self.scene.num_worlds, self.scene.bodies_per_world
        off = np.ascontiguousarray(world_offset, np.float32)
        rows = np.ascontiguousarray(view_rows, np.int32)
        cam = np.ascontiguousarray(camera_offset, np.float32)
        ys = np.ascontiguousarray(y_scale, np.float32)
        V = rows.shape[1]
        tr = np.zeros((W, N, 12), np.float32)
        views = np.zeros((W, V, 8), np.float32)
        self.lib.orc_feeder(self.h, off.ctypes.data, tr.ctypes.data, V, rows.ctypes.data, cam.ctypes.data,
                            ys.ctypes.data, views.ctypes.data)
        return tr, views

    def indeterminate(self):
        out = np.zeros(self.scene.num_worlds, np.int64)
        self.lib.orc_indeterminate(self.h, out.ctypes.data)
        return out

    def get_hull(self, obj):
        he = np.zeros((256, 4), np.uint32)
        ed = np.zeros(256, np.uint32)
        po = np.zeros(256, np.uint32)
        pl = np.zeros((256, 4), np.float32)
        cnt = np.zeros(4, np.int32)
        self.lib.orc_get_hull(self.h, obj, he.ctypes.data, ed.ctypes.data, po.ctypes.data,
                              pl.ctypes.data, cnt.ctypes.data)
        return dict(half_edges=he[:cnt[0]].copy(), edges=ed[:cnt[1]].copy(),
                    polys=po[:cnt[2]].copy(), planes=pl[:cnt[2]].copy(), counts=cnt)
