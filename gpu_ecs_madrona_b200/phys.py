"""Python host mirror of the rigid-body world-step executor.

Thin ctypes binding of include/madrona_b200.h (libmadrona_b200.so, CUDA,
sm_100a). The class mirrors how the reference is driven for this path:
construct an executor from a config + shared collision objects + per-world
init data (MWCudaExecutor(StateConfig, CompileConfig), include/madrona/
mw_gpu.hpp:20-76 with RigidBodyPhysicsSystem as the world's systems,
include/madrona/physics.hpp:419-447), call ``run()`` per step
(src/mw/cuda_exec.cpp:1777-1794) and read exported columns in place
(``getExported``). There is no CPU path: importing works anywhere, creating an
executor without the CUDA library or without a GPU raises.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# MDR_LIB selects another build of the same library (e.g. the FMA-contracted
# tolerance build libmadrona_b200_fma.so); the default is the bit-exact build.
LIB_PATH = os.environ.get("MDR_LIB") or os.path.join(HERE, "libmadrona_b200.so")
if not os.path.isabs(LIB_PATH):
    LIB_PATH = os.path.join(HERE, LIB_PATH)

COL_BYTES = [8, 12, 16, 12, 24, 4, 4, 28, 28, 24, 12, 12, 4]
(COL_ENTITY, COL_POSITION, COL_ROTATION, COL_SCALE, COL_VELOCITY, COL_OBJECT_ID,
 COL_RESPONSE_TYPE, COL_SUBSTEP_PREV, COL_PRESOLVE_POS, COL_PRESOLVE_VEL,
 COL_EXT_FORCE, COL_EXT_TORQUE, COL_LEAF_ID) = range(13)

MDR_OK, MDR_EINVAL, MDR_ENODEVICE, MDR_ECAPACITY, MDR_EOVERFLOW, MDR_ESTATE = 0, -1, -2, -3, -4, -5
_ERR = {-1: "invalid argument", -2: "no CUDA device", -3: "capacity does not fit the device",
        -4: "per-world capacity overflowed", -5: "call in the wrong phase"}


class MdrObject(C.Structure):
    _fields_ = [("type", C.c_int32), ("inv_inertia", C.c_float * 3),
                ("inv_mass", C.c_float), ("mu_s", C.c_float), ("mu_d", C.c_float),
                ("aabb_min", C.c_float * 3), ("aabb_max", C.c_float * 3),
                ("vert_offset", C.c_int32), ("num_verts", C.c_int32),
                ("poly_offset", C.c_int32), ("num_poly_words", C.c_int32)]


class MdrPhysConfig(C.Structure):
    _fields_ = [("num_worlds", C.c_int32), ("max_bodies_per_world", C.c_int32),
                ("num_substeps", C.c_int32), ("max_contacts", C.c_int32),
                ("delta_t", C.c_float), ("gravity", C.c_float * 3),
                ("gpu_id", C.c_int32), ("block_threads", C.c_int32),
                ("smem_contacts", C.c_int32), ("worlds_per_block", C.c_int32)]


CONTACT_DTYPE = np.dtype([("ref_row", "<i4"), ("ref_pad", "<i4"), ("alt_row", "<i4"),
                          ("alt_pad", "<i4"), ("points", "<f4", (4, 4)),
                          ("num_points", "<i4"), ("normal", "<f4", (3,)),
                          ("lambda_n", "<f4", (4,))])
assert CONTACT_DTYPE.itemsize == 112


class MadronaError(RuntimeError):
    def __init__(self, code, what):
        super().__init__(f"{what}: {_ERR.get(code, code)} ({code})")
        self.code = code


def bind(lib):
    """Declares the C ABI on a loaded library."""
    vp, i32 = C.c_void_p, C.c_int32
    lib.mdr_phys_create.argtypes = [vp, vp, i32, vp, vp, C.POINTER(vp)]
    lib.mdr_phys_destroy.argtypes = [vp]
    lib.mdr_phys_destroy.restype = None
    lib.mdr_phys_set_row_counts.argtypes = [vp, vp]
    lib.mdr_phys_write_column.argtypes = [vp, i32, vp]
    lib.mdr_phys_read_column.argtypes = [vp, i32, vp]
    lib.mdr_phys_flag_rebuild.argtypes = [vp]
    if hasattr(lib, "mdr_phys_reset_worlds"):
        lib.mdr_phys_reset_worlds.argtypes = [vp, vp, i32]
        lib.mdr_phys_write_world_column.argtypes = [vp, i32, i32, i32, vp]
        lib.mdr_phys_read_world_column.argtypes = [vp, i32, i32, i32, vp]
        lib.mdr_phys_set_world_rows.argtypes = [vp, i32, i32]
    lib.mdr_phys_set_joints.argtypes = [vp, i32, vp, vp]
    lib.mdr_phys_step.argtypes = [vp, i32]
    lib.mdr_phys_step_async.argtypes = [vp, i32]
    lib.mdr_phys_sync.argtypes = [vp]
    lib.mdr_phys_probe_step.argtypes = [vp, i32, i32, vp, vp, vp, vp]
    lib.mdr_phys_stats.argtypes = [vp, vp, vp]
    lib.mdr_phys_read_bvh.argtypes = [vp, vp, vp, vp]
    lib.mdr_phys_get_hull.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    lib.mdr_version.restype = C.c_char_p
    if hasattr(lib, "mdr_load_hull_obj"):
        lib.mdr_load_hull_obj.argtypes = [C.c_char_p, C.c_int64, vp, vp, vp, vp, vp, C.c_char_p, i32]
        lib.mdr_load_hull_obj_file.argtypes = [C.c_char_p, vp, vp, vp, vp, vp, C.c_char_p, i32]
    if hasattr(lib, "mdr_phys_pack_instances"):
        lib.mdr_phys_pack_instances.argtypes = [vp, vp, vp, vp]
        lib.mdr_phys_pack_views.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    if hasattr(lib, "mdr_phys_trace_rays"):
        lib.mdr_phys_trace_rays.argtypes = [vp, i32] + [vp] * 7
        lib.mdr_phys_trace_rays_device.argtypes = [vp, i32] + [vp] * 7
    for name in ("mdr_phys_column_ptr", "mdr_phys_stream"):
        if hasattr(lib, name):
            getattr(lib, name).restype = vp
    if hasattr(lib, "mdr_phys_column_ptr"):
        lib.mdr_phys_column_ptr.argtypes = [vp, i32]
        lib.mdr_phys_stream.argtypes = [vp]
        lib.mdr_phys_launch_info.argtypes = [vp, vp]
        lib.mdr_phys_kernel_ms.argtypes = [vp, vp]
        lib.mdr_phys_kernel_ms.restype = C.c_float
        if hasattr(lib, "mdr_phys_last_launches"):
            lib.mdr_phys_last_launches.argtypes = [vp]
        lib.mdr_phys_row_stride.argtypes = [vp]
        lib.mdr_phys_write_column_raw_async.argtypes = [vp, i32, vp]
        lib.mdr_phys_read_column_raw_async.argtypes = [vp, i32, vp]
        if hasattr(lib, "mdr_phys_step_io_async"):
            lib.mdr_phys_step_io_async.argtypes = [vp, i32, vp, vp, i32, vp, vp, i32]
        lib.mdr_host_alloc.argtypes = [C.c_uint64]
        lib.mdr_host_alloc.restype = vp
        lib.mdr_host_free.argtypes = [vp]
        lib.mdr_host_free.restype = None
    return lib


_LIB = None


def load_library():
    """Loads the CUDA library; fails loudly when it has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a). There is no CPU fallback.")
        _LIB = bind(C.CDLL(LIB_PATH))
    return _LIB


def pack_objects(objects):
    """ObjectDef list (scenes.py) -> (mdr_object array, verts [V,3] f32, polygon words u32)."""
    verts, words, arr = [], [], (MdrObject * len(objects))()
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
    v = np.concatenate(verts) if voff else np.zeros((1, 3), np.float32)
    w = np.concatenate(words) if woff else np.zeros(1, np.uint32)
    return arr, np.ascontiguousarray(v, np.float32), np.ascontiguousarray(w, np.uint32)


class PhysExecutor:
    """Batched rigid-body worlds on one GPU. ``scene`` is a ``scenes.Scene``."""

    def __init__(self, scene, gpu_id=0, block_threads=0, smem_contacts=0, worlds_per_block=0, lib=None):
        self.gpu_id = gpu_id
        self.lib = lib if lib is not None else load_library()
        self.scene = scene
        self.W, self.N = scene.num_worlds, scene.bodies_per_world
        cfg = MdrPhysConfig(scene.num_worlds, scene.bodies_per_world, scene.num_substeps,
                            scene.max_contacts, scene.delta_t, (C.c_float * 3)(*scene.gravity),
                            gpu_id, block_threads, smem_contacts, worlds_per_block)
        objs, verts, words = pack_objects(scene.objects)
        h = C.c_void_p()
        rc = self.lib.mdr_phys_create(C.byref(cfg), objs, len(scene.objects),
                                      verts.ctypes.data, words.ctypes.data, C.byref(h))
        if rc != MDR_OK:
            raise MadronaError(rc, "mdr_phys_create")
        self.h = h
        rows = getattr(scene, "rows_per_world", None)
        if rows is None:
            rows = np.full(self.W, self.N, np.int32)
        self.set_row_counts(rows)
        vel = np.ascontiguousarray(scene.vel, np.float32)
        for col, arr, dt in ((COL_POSITION, scene.pos, np.float32), (COL_ROTATION, scene.rot, np.float32),
                             (COL_SCALE, scene.scale, np.float32), (COL_VELOCITY, vel, np.float32),
                             (COL_OBJECT_ID, scene.obj_ids, np.int32),
                             (COL_RESPONSE_TYPE, scene.response, np.uint32),
                             (COL_EXT_FORCE, scene.ext_force, np.float32),
                             (COL_EXT_TORQUE, scene.ext_torque, np.float32)):
            self.write(col, np.ascontiguousarray(arr, dt))
        if getattr(scene, "joints", None) is not None:
            self.set_joints(scene.joints, scene.joint_counts)

    def set_joints(self, joints, counts=None):
        """Replaces the joint constraints: ``joints`` [W, J] of ``scenes.JOINT_DTYPE`` (the
        ConstraintData rows of each world, solved in this order after the contacts of every
        substep), ``counts`` [W] live joints per world (default: all J). ``None`` removes them."""
        if joints is None:
            self._check(self.lib.mdr_phys_set_joints(self.h, 0, None, None), "set_joints")
            return
        from .scenes import JOINT_DTYPE
        joints = np.ascontiguousarray(joints, JOINT_DTYPE)
        assert joints.ndim == 2 and joints.shape[0] == self.W
        if counts is None:
            counts = np.full(self.W, joints.shape[1], np.int32)
        counts = np.ascontiguousarray(counts, np.int32)
        self._check(self.lib.mdr_phys_set_joints(self.h, joints.shape[1], counts.ctypes.data,
                                                 joints.ctypes.data), "set_joints")

    def _check(self, rc, what):
        if rc != MDR_OK:
            raise MadronaError(rc, what)

    def close(self):
        if getattr(self, "h", None):
            self.lib.mdr_phys_destroy(self.h)
            self.h = None
            for ptr in getattr(self, "_pinned", []):
                self.lib.mdr_host_free(ptr)
            self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_row_counts(self, rows):
        rows = np.ascontiguousarray(rows, np.int32)
        assert rows.shape == (self.W,)
        self._check(self.lib.mdr_phys_set_row_counts(self.h, rows.ctypes.data), "set_row_counts")

    def write(self, col, arr):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes == self.W * self.N * COL_BYTES[col], (col, arr.shape)
        self._check(self.lib.mdr_phys_write_column(self.h, col, arr.ctypes.data), "write_column")

    def read(self, col):
        out = np.empty((self.W, self.N, COL_BYTES[col]), np.uint8)
        self._check(self.lib.mdr_phys_read_column(self.h, col, out.ctypes.data), "read_column")
        return out

    def read_f32(self, col):
        return self.read(col).view(np.float32)

    def state(self):
        """(pos[W,N,3], rot[W,N,4], vel[W,N,6])."""
        return (self.read_f32(COL_POSITION), self.read_f32(COL_ROTATION),
                self.read_f32(COL_VELOCITY))

    def run(self, num_steps=1):
        """MWCudaExecutor::run(): blocking."""
        self._check(self.lib.mdr_phys_step(self.h, num_steps), "step")

    step = run

    def run_async(self, num_steps=1):
        self._check(self.lib.mdr_phys_step_async(self.h, num_steps), "step_async")

    def sync(self):
        self._check(self.lib.mdr_phys_sync(self.h), "sync")

    def flag_rebuild(self):
        self._check(self.lib.mdr_phys_flag_rebuild(self.h), "flag_rebuild")

    def reset_worlds(self, worlds):
        """RigidBodyPhysicsSystem::reset of the listed worlds: only their BVHs are rebuilt before the next step."""
        w = np.ascontiguousarray(worlds, np.int32)
        self._check(self.lib.mdr_phys_reset_worlds(self.h, w.ctypes.data, len(w)), "reset_worlds")

    def write_world(self, col, world, arr):
        """One world's rows of a column ([rows, ...] of the column's element type)."""
        a = np.ascontiguousarray(arr)
        rows = a.nbytes // COL_BYTES[col]
        self._check(self.lib.mdr_phys_write_world_column(self.h, col, world, rows, a.ctypes.data), "write_world_column")

    def read_world(self, col, world, rows=None):
        rows = self.N if rows is None else rows
        out = np.empty((rows, COL_BYTES[col]), np.uint8)
        self._check(self.lib.mdr_phys_read_world_column(self.h, col, world, rows, out.ctypes.data), "read_world_column")
        return out

    def probe_step(self, world, max_pairs=4096):
        S, MC = self.scene.num_substeps, self.scene.max_contacts
        npairs = np.zeros(1, np.int32)
        pairs = np.zeros((max_pairs, 2), np.int32)
        ncon = np.zeros(S, np.int32)
        contacts = np.zeros((S, MC), CONTACT_DTYPE)
        self._check(self.lib.mdr_phys_probe_step(self.h, world, max_pairs, npairs.ctypes.data,
                                                 pairs.ctypes.data, ncon.ctypes.data,
                                                 contacts.ctypes.data), "probe_step")
        return dict(pairs=pairs[:npairs[0]].copy(), num_contacts=ncon,
                    contacts=[contacts[s, :ncon[s]].copy() for s in range(S)])

    def stats(self):
        p = np.zeros(self.W, np.int32)
        c = np.zeros(self.W, np.int32)
        self._check(self.lib.mdr_phys_stats(self.h, p.ctypes.data, c.ctypes.data), "stats")
        return p, c

    def read_bvh(self):
        slot = np.zeros((self.W, self.N, 6), np.float32)
        leaf = np.zeros((self.W, self.N, 6), np.float32)
        order = np.zeros((self.W, self.N), np.uint16)
        self._check(self.lib.mdr_phys_read_bvh(self.h, slot.ctypes.data, leaf.ctypes.data,
                                               order.ctypes.data), "read_bvh")
        return slot, leaf, order

    def get_hull(self, obj):
        he = np.zeros((1024, 4), np.uint32)
        ed = np.zeros(1024, np.uint32)
        po = np.zeros(1024, np.uint32)
        pl = np.zeros((1024, 4), np.float32)
        cnt = np.zeros(4, np.int32)
        self._check(self.lib.mdr_phys_get_hull(self.h, obj, he.ctypes.data, ed.ctypes.data,
                                               po.ctypes.data, pl.ctypes.data, cnt.ctypes.data),
                    "get_hull")
        return dict(half_edges=he[:cnt[0]].copy(), edges=ed[:cnt[1]].copy(),
                    polys=po[:cnt[2]].copy(), planes=pl[:cnt[2]].copy(), counts=cnt)

    # ---- CUDA-only entry points -------------------------------------------
    def column_ptr(self, col):
        return self.lib.mdr_phys_column_ptr(self.h, col)

    def column_tensor(self, col, dtype=np.float32):
        """Zero-copy torch view [num_worlds, row_stride, bytes / itemsize] of an exported column
        (getExported + Tensor.to_torch of the reference, include/madrona/python.hpp:21-73)."""
        from .tensor import as_torch
        shape = (self.W, self.row_stride(), COL_BYTES[col] // np.dtype(dtype).itemsize)
        return as_torch(self.column_ptr(col), shape, dtype, self.gpu_id, owner=self)

    def tensor(self, col, dtype=np.float32):
        """madrona_python.Tensor of an exported column (what Executor::getExported is wrapped in)."""
        from .tensor import as_tensor
        shape = (self.W, self.row_stride(), COL_BYTES[col] // np.dtype(dtype).itemsize)
        return as_tensor(self.column_ptr(col), shape, dtype, self.gpu_id)

    def cuda_sync(self):
        """madrona_python.CudaSync on the executor's stream: .wait(torch.cuda.current_stream().cuda_stream)
        orders a consumer's stream after the steps enqueued so far (run_async), without a host sync."""
        from .tensor import cuda_sync
        self.lib.mdr_phys_done_event.restype = C.c_void_p
        self.lib.mdr_phys_done_event.argtypes = [C.c_void_p]
        return cuda_sync(self.lib.mdr_phys_done_event(self.h))

    def row_stride(self):
        return int(self.lib.mdr_phys_row_stride(self.h))

    def pinned(self, col):
        """A page-locked host staging array [W, row_stride, bytes] for column `col`."""
        shape = (self.W, self.row_stride(), COL_BYTES[col])
        n = int(np.prod(shape))
        ptr = self.lib.mdr_host_alloc(n)
        if not ptr:
            raise MemoryError("mdr_host_alloc")
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(ptr)
        return np.ctypeslib.as_array((C.c_uint8 * n).from_address(ptr)).reshape(shape)

    def write_raw_async(self, col, arr):
        self._check(self.lib.mdr_phys_write_column_raw_async(self.h, col, arr.ctypes.data), "write_raw")

    def read_raw_async(self, col, arr):
        self._check(self.lib.mdr_phys_read_column_raw_async(self.h, col, arr.ctypes.data), "read_raw")

    def step_io_async(self, ins, outs, chunks=4):
        """One step with host-side inputs / outputs [(column, raw host array)], pipelined
        over `chunks` world ranges (copies overlap the step). sync() waits for it."""
        def pack(pairs):
            cols = (C.c_int32 * len(pairs))(*[c for c, _ in pairs])
            ptrs = (C.c_void_p * len(pairs))(*[a.ctypes.data for _, a in pairs])
            return cols, ptrs
        ic, ip = pack(ins)
        oc, op = pack(outs)
        self._check(self.lib.mdr_phys_step_io_async(self.h, len(ins), ic, ip, len(outs), oc, op, chunks),
                    "step_io")

    def pack_instances(self, out_transforms, out_object_ids, world_offset=None):
        """instanceTransformSetup into buffers [W, row_stride, 12] f32 / [W, row_stride] i32
        given as raw addresses (device pointers on the CUDA library). Enqueues only."""
        off = None if world_offset is None else np.ascontiguousarray(world_offset, np.float32)
        self._check(self.lib.mdr_phys_pack_instances(self.h, None if off is None else off.ctypes.data,
                                                     out_transforms, out_object_ids), "pack_instances")

    def pack_views(self, view_rows, camera_offset, y_scale, out_views, views_per_world, world_offset=None):
        """updateViewData into [W, V, 8] f32; all arrays raw addresses (device pointers on CUDA)."""
        off = None if world_offset is None else np.ascontiguousarray(world_offset, np.float32)
        self._check(self.lib.mdr_phys_pack_views(self.h, views_per_world, view_rows, camera_offset, y_scale,
                                                 None if off is None else off.ctypes.data, out_views),
                    "pack_views")

    def trace_rays(self, origins, dirs, t_max=None):
        """BVH::traceRay for [W, R, 3] host rays: (t[W,R], normal[W,R,3], entity[W,R,2], row[W,R])."""
        o = np.ascontiguousarray(origins, np.float32)
        d = np.ascontiguousarray(dirs, np.float32)
        assert o.shape == d.shape and o.shape[0] == self.W and o.shape[2] == 3
        R = o.shape[1]
        tm = None if t_max is None else np.ascontiguousarray(np.broadcast_to(np.asarray(t_max, np.float32), (self.W, R)))
        t = np.zeros((self.W, R), np.float32)
        nrm = np.zeros((self.W, R, 3), np.float32)
        ent = np.zeros((self.W, R, 2), np.int32)
        row = np.zeros((self.W, R), np.int32)
        self._check(self.lib.mdr_phys_trace_rays(self.h, R, o.ctypes.data, d.ctypes.data,
                                                 None if tm is None else tm.ctypes.data, t.ctypes.data,
                                                 nrm.ctypes.data, ent.ctypes.data, row.ctypes.data), "trace_rays")
        return t, nrm, ent, row

    def launch_info(self):
        info = np.zeros(6, np.int32)
        self._check(self.lib.mdr_phys_launch_info(self.h, info.ctypes.data), "launch_info")
        return dict(blocks=int(info[0]), threads=int(info[1]), smem_bytes=int(info[2]),
                    blocks_per_sm=int(info[3]), threads_per_world=int(info[4]),
                    worlds_per_block=int(info[5]))

    def kernel_ms(self):
        """(ms per step, steps) timed on the device since the last call."""
        n = C.c_int32(0)
        ms = self.lib.mdr_phys_kernel_ms(self.h, C.byref(n))
        return float(ms), int(n.value)

    def last_launches(self):
        """Step kernels launched by the last run / run_async call (a call of n > 1 steps is one batch launch)."""
        return int(self.lib.mdr_phys_last_launches(self.h))
