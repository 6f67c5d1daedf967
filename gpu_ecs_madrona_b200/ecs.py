"""Python host mirror of the ECS state layer + task-graph executor.

Thin ctypes binding of the mdr_ecs_* part of include/madrona_b200.h. `EcsRegistry`
mirrors the reference's registration / graph-building calls (ECSRegistry,
include/madrona/state.hpp:109-134; TaskGraph::Builder, include/madrona/taskgraph.hpp:
15-97) for worlds assembled from Python; `EcsExecutor` is MWCudaExecutor for one of
the built-in worlds (include/madrona/mw_gpu.hpp:55-76): construct, `run()`, read
exported columns. There is no CPU path: creating an executor without the CUDA library
or without a GPU raises.
"""
import ctypes as C

import numpy as np

from . import ecs_apps
from .phys import MadronaError, load_library, MDR_OK

OP_DTYPE = np.dtype([("op", "<i4"), ("a", "<i4"), ("b", "<i4")])
ARCH_FIXED = 1
COMP_ENTITY = 0


class MdrEcsConfig(C.Structure):
    _fields_ = [("num_worlds", C.c_int32), ("first_world", C.c_int32), ("gpu_id", C.c_int32),
                ("ids_per_world", C.c_int32), ("max_script_handles", C.c_int32), ("app", C.c_int32)]


_bound = False


def _lib():
    global _bound
    lib = load_library()
    if not _bound:
        vp, i32 = C.c_void_p, C.c_int32
        lib.mdr_ecs_create.argtypes = [vp, C.POINTER(vp)]
        lib.mdr_ecs_destroy.argtypes = [vp]
        lib.mdr_ecs_destroy.restype = None
        lib.mdr_ecs_register_component.argtypes = [vp, i32, i32]
        lib.mdr_ecs_register_archetype.argtypes = [vp, vp, i32, i32, i32]
        lib.mdr_ecs_register_singleton.argtypes = [vp, i32, i32]
        lib.mdr_ecs_export_column.argtypes = [vp, i32, i32, i32]
        lib.mdr_ecs_add_parallel_for.argtypes = [vp, i32, vp, i32, vp, i32]
        lib.mdr_ecs_add_clear_tmp.argtypes = [vp, i32, vp, i32]
        lib.mdr_ecs_add_reset_tmp_alloc.argtypes = [vp, vp, i32]
        lib.mdr_ecs_add_sort.argtypes = [vp, i32, i32, i32, vp, i32]
        lib.mdr_ecs_build.argtypes = [vp]
        lib.mdr_ecs_make_entities.argtypes = [vp, i32, i32]
        lib.mdr_ecs_script.argtypes = [vp, vp, vp]
        lib.mdr_ecs_write_column.argtypes = [vp, i32, i32, i32, vp]
        lib.mdr_ecs_read_column.argtypes = [vp, i32, i32, i32, vp]
        lib.mdr_ecs_row_counts.argtypes = [vp, i32, vp]
        for name in ("mdr_ecs_num_archetypes", "mdr_ecs_id_capacity", "mdr_ecs_build", "mdr_ecs_sync",
                     "mdr_ecs_launches_per_step", "mdr_ecs_last_launches"):
            getattr(lib, name).argtypes = [vp]
        lib.mdr_ecs_num_columns.argtypes = [vp, i32]
        lib.mdr_ecs_column_bytes.argtypes = [vp, i32, i32]
        lib.mdr_ecs_row_stride.argtypes = [vp, i32]
        lib.mdr_ecs_column_ptr.argtypes = [vp, i32, i32]
        lib.mdr_ecs_column_ptr.restype = vp
        lib.mdr_ecs_read_idmap.argtypes = [vp, i32, vp, vp, vp, vp]
        lib.mdr_ecs_step.argtypes = [vp, i32]
        lib.mdr_ecs_step_async.argtypes = [vp, i32]
        lib.mdr_ecs_stream.argtypes = [vp]
        lib.mdr_ecs_stream.restype = vp
        lib.mdr_ecs_step_ms.argtypes = [vp, vp]
        lib.mdr_ecs_step_ms.restype = C.c_float
        lib.mdr_ecs_get_exported.argtypes = [vp, i32]
        lib.mdr_static_map_build.argtypes = [vp, vp, i32, vp, vp]
        lib.mdr_static_map_lookup.argtypes = [vp, vp, C.c_uint32]
        lib.mdr_ecs_column_map.argtypes = [vp, i32, vp, vp]
        lib.mdr_ecs_lookup_columns.argtypes = [vp, i32, vp, i32, vp]
        lib.mdr_ecs_get_exported.restype = vp
        lib.mdr_ecs_export_rows.argtypes = [vp, i32]
        lib.mdr_ecs_export_rows.restype = C.c_int64
        lib.mdr_ecs_export_bytes_per_row.argtypes = [vp, i32]
        lib.mdr_ecs_read_exported.argtypes = [vp, i32, vp, C.c_int64]
        lib.mdr_ecs_read_exported.restype = C.c_int64
        lib.mdr_app_setup.argtypes = [vp, i32, vp, i32]
        _bound = True
    return lib


def _check(rc, what):
    if rc < 0:
        raise MadronaError(rc, what)
    return rc


def _ids(values):
    a = np.asarray(list(values), np.int32)
    return a, (a.ctypes.data if len(a) else None), len(a)


class EcsRegistry:
    """Registration + graph building + execution over the C ABI, for worlds assembled by hand
    (the systems still have to exist in the compiled dispatch table, csrc/ecs_apps.cuh)."""

    def __init__(self, num_worlds, gpu_id=0, first_world=0, ids_per_world=0, max_script_handles=0):
        self.lib = _lib()
        self.num_worlds = num_worlds
        self.gpu_id = gpu_id
        cfg = MdrEcsConfig(num_worlds, first_world, gpu_id, ids_per_world, max_script_handles, 0)
        h = C.c_void_p()
        _check(self.lib.mdr_ecs_create(C.byref(cfg), C.byref(h)), "mdr_ecs_create")
        self.h = h

    @classmethod
    def attach(cls, handle, num_worlds, gpu_id=0):
        """View of an executor somebody else owns (an out-of-tree world's TaskGraphExecutor, examples/):
        the state accessors of this class on its mdr_ecs handle; close() leaves it alone."""
        self = cls.__new__(cls)
        self.lib = _lib()
        self.num_worlds, self.gpu_id = num_worlds, gpu_id
        self.h = C.c_void_p(handle)
        self._borrowed = True
        self._describe()
        return self

    def close(self):
        if self.h and not getattr(self, "_borrowed", False):
            self.lib.mdr_ecs_destroy(self.h)
        self.h = None
        for ptr in getattr(self, "_pinned", []):
            self.lib.mdr_host_free.argtypes = [C.c_void_p]
            self.lib.mdr_host_free(ptr)
        self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- ECSRegistry ----
    def register_component(self, nbytes, align=4):
        return _check(self.lib.mdr_ecs_register_component(self.h, nbytes, align), "register_component")

    def register_archetype(self, comps, max_rows, fixed=False):
        a, p, n = _ids(comps)
        return _check(self.lib.mdr_ecs_register_archetype(self.h, p, n, max_rows, ARCH_FIXED if fixed else 0),
                      "register_archetype")

    def register_singleton(self, nbytes, align=4):
        return _check(self.lib.mdr_ecs_register_singleton(self.h, nbytes, align), "register_singleton")

    def export_column(self, arch, comp, slot):
        _check(self.lib.mdr_ecs_export_column(self.h, arch, comp, slot), "export_column")

    # ---- TaskGraph::Builder ----
    def parallel_for(self, func, comps, deps=()):
        a, p, n = _ids(comps)
        d, dp, dn = _ids(deps)
        return _check(self.lib.mdr_ecs_add_parallel_for(self.h, func, p, n, dp, dn), "parallel_for")

    def clear_tmp(self, arch, deps=()):
        d, dp, dn = _ids(deps)
        return _check(self.lib.mdr_ecs_add_clear_tmp(self.h, arch, dp, dn), "clear_tmp")

    def reset_tmp_alloc(self, deps=()):
        d, dp, dn = _ids(deps)
        return _check(self.lib.mdr_ecs_add_reset_tmp_alloc(self.h, dp, dn), "reset_tmp_alloc")

    def sort_archetype(self, arch, key_comp, remap_entities=False, deps=()):
        d, dp, dn = _ids(deps)
        return _check(self.lib.mdr_ecs_add_sort(self.h, arch, key_comp, int(remap_entities), dp, dn), "sort")

    def build(self):
        _check(self.lib.mdr_ecs_build(self.h), "build")
        self._describe()

    def _describe(self):
        self.num_archetypes = self.lib.mdr_ecs_num_archetypes(self.h)
        self.col_bytes = [[self.lib.mdr_ecs_column_bytes(self.h, a, c)
                           for c in range(self.lib.mdr_ecs_num_columns(self.h, a))]
                          for a in range(self.num_archetypes)]
        self.stride = [self.lib.mdr_ecs_row_stride(self.h, a) for a in range(self.num_archetypes)]
        self.id_cap = self.lib.mdr_ecs_id_capacity(self.h)

    # ---- world constructors / state access ----
    def make_entities(self, arch, count):
        _check(self.lib.mdr_ecs_make_entities(self.h, arch, count), "make_entities")

    def script(self, per_world_ops):
        """per_world_ops: one OP_DTYPE array per world (immediate mode, in order)."""
        offs = np.zeros(self.num_worlds + 1, np.int32)
        offs[1:] = np.cumsum([len(o) for o in per_world_ops])
        if offs[-1] == 0:
            return
        ops = np.ascontiguousarray(np.concatenate([np.asarray(o, OP_DTYPE) for o in per_world_ops]))
        _check(self.lib.mdr_ecs_script(self.h, ops.ctypes.data, offs.ctypes.data), "script")

    def row_counts(self):
        out = np.zeros((self.num_archetypes, self.num_worlds), np.int32)
        for a in range(self.num_archetypes):
            _check(self.lib.mdr_ecs_row_counts(self.h, a, out[a].ctypes.data), "row_counts")
        return np.ascontiguousarray(out.T)

    def read_column_all(self, arch, col, rows=None):
        """[num_worlds, rows, bytes] u8 of one column (rows defaults to the row stride)."""
        rows = self.stride[arch] if rows is None else rows
        out = np.zeros((self.num_worlds, rows, self.col_bytes[arch][col]), np.uint8)
        if rows:
            _check(self.lib.mdr_ecs_read_column(self.h, arch, col, rows, out.ctypes.data), "read_column")
        return out

    def write_column_all(self, arch, col, data):
        """data: [num_worlds, rows, ...] of the column's element type."""
        d = np.ascontiguousarray(data)
        rows = d.shape[1]
        d = d.view(np.uint8).reshape(self.num_worlds, rows, self.col_bytes[arch][col])
        _check(self.lib.mdr_ecs_write_column(self.h, arch, col, rows, d.ctypes.data), "write_column")

    def idmap(self, world):
        a, b = np.zeros(self.id_cap, np.int32), np.zeros(self.id_cap, np.int32)
        gen, state = np.zeros(self.id_cap, np.uint32), np.zeros(10, np.int32)
        n = _check(self.lib.mdr_ecs_read_idmap(self.h, world, a.ctypes.data, b.ctypes.data, gen.ctypes.data,
                                               state.ctypes.data), "read_idmap")
        return a[:n], b[:n], gen[:n], state

    def column_ptr(self, arch, col):
        return self.lib.mdr_ecs_column_ptr(self.h, arch, col)

    # ---- executor ----
    def run(self, num_steps=1):
        """MWCudaExecutor::run(), num_steps times (blocking)."""
        _check(self.lib.mdr_ecs_step(self.h, num_steps), "step")

    def run_async(self, num_steps=1):
        _check(self.lib.mdr_ecs_step_async(self.h, num_steps), "step_async")

    def sync(self):
        _check(self.lib.mdr_ecs_sync(self.h), "sync")

    def step_ms(self):
        n = C.c_int32(0)
        ms = self.lib.mdr_ecs_step_ms(self.h, C.byref(n))
        return ms, n.value

    def launches_per_step(self):
        return self.lib.mdr_ecs_launches_per_step(self.h)

    def last_launches(self):
        """Kernels launched by the last run / run_async call (a call of n > 1 steps is one batch launch)."""
        return self.lib.mdr_ecs_last_launches(self.h)

    def column_map(self, arch):
        """The archetype's perfect-hash column map (ColumnMap columnLookup of the reference): (keys[128], values[128])."""
        k, v = np.zeros(128, np.uint32), np.zeros(128, np.uint32)
        _check(self.lib.mdr_ecs_column_map(self.h, arch, k.ctypes.data, v.ctypes.data), "mdr_ecs_column_map")
        return k, v

    def lookup_columns(self, arch, comp_ids):
        """Column index of every component id in `arch` (-1: absent), resolved ON THE DEVICE through the
        uploaded map, as Context::get<T>(Loc) does."""
        ids = np.ascontiguousarray(comp_ids, np.int32)
        out = np.zeros(len(ids), np.int32)
        _check(self.lib.mdr_ecs_lookup_columns(self.h, arch, ids.ctypes.data, len(ids), out.ctypes.data),
               "mdr_ecs_lookup_columns")
        return out

    def get_exported(self, slot):
        """(device pointer, rows, bytes per row) of an exported column: getExported(slot)."""
        return (self.lib.mdr_ecs_get_exported(self.h, slot), self.lib.mdr_ecs_export_rows(self.h, slot),
                self.lib.mdr_ecs_export_bytes_per_row(self.h, slot))

    def export_tensor(self, slot, dtype=np.float32):
        """Zero-copy torch view of an exported column, [rows, bytes_per_row / itemsize]: what a trainer
        holds (Tensor.to_torch() of the reference). Staged exports change length with the row counts:
        take a new view after a step that created or destroyed rows."""
        from .tensor import as_torch
        ptr, rows, nb = self.get_exported(slot)
        return as_torch(ptr, (rows, nb // np.dtype(dtype).itemsize), dtype, self.gpu_id, owner=self)

    def column_tensor(self, arch, col, dtype=np.float32):
        """Zero-copy torch view of a whole table column, [num_worlds, row_stride, bytes / itemsize]."""
        from .tensor import as_torch
        nb = self.col_bytes[arch][col]
        return as_torch(self.column_ptr(arch, col), (self.num_worlds, self.stride[arch], nb // np.dtype(dtype).itemsize),
                        dtype, self.gpu_id, owner=self)

    def pinned_export_buffer(self, slot):
        """A page-locked host array large enough for any row count of exported column `slot` (u8)."""
        a = self.lib.mdr_ecs_get_exported  # noqa: F841  (the library is loaded)
        arch_rows = max(self.stride) * self.num_worlds
        nb = self.lib.mdr_ecs_export_bytes_per_row(self.h, slot)
        n = int(arch_rows) * int(nb)
        self.lib.mdr_host_alloc.restype = C.c_void_p
        self.lib.mdr_host_alloc.argtypes = [C.c_uint64]
        ptr = self.lib.mdr_host_alloc(n)
        if not ptr:
            raise MemoryError("mdr_host_alloc")
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(ptr)
        return np.ctypeslib.as_array((C.c_uint8 * n).from_address(ptr))

    def export_into(self, slot, host):
        """Copies exported column `slot` into the (pinned) host array; returns the bytes copied."""
        n = self.lib.mdr_ecs_read_exported(self.h, slot, host.ctypes.data, host.nbytes)
        _check(n, "read_exported")
        return int(n)

    def step_read(self, slots, hosts):
        """One step, then exported columns `slots` into the (pinned) host arrays `hosts`, in one call with two host
        synchronisations (mdr_ecs_step_read); returns the bytes copied per slot."""
        n = len(slots)
        lib = self.lib
        lib.mdr_ecs_step_read.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        sl = (C.c_int32 * n)(*slots)
        ptrs = (C.c_void_p * n)(*[h.ctypes.data for h in hosts])
        mx = (C.c_int64 * n)(*[h.nbytes for h in hosts])
        out = (C.c_int64 * n)()
        _check(lib.mdr_ecs_step_read(self.h, n, sl, ptrs, mx, out), "step_read")
        return [int(v) for v in out]

    def export(self, slot):
        """Host copy of an exported column (u8)."""
        ptr, rows, nb = self.get_exported(slot)
        out = np.zeros(max(rows * nb, 1), np.uint8)
        n = self.lib.mdr_ecs_read_exported(self.h, slot, out.ctypes.data, out.nbytes)
        _check(n, "read_exported")
        return out[:n].copy()


class EcsExecutor(EcsRegistry):
    """One of the built-in task-graph worlds (ecs_apps.APP_*), constructed like the reference's
    MWCudaExecutor: registerTypes + setupTasks + per-world constructors, then initial columns."""

    def __init__(self, app, num_worlds, params=None, gpu_id=0, first_world=0, init=True):
        super().__init__(num_worlds, gpu_id, first_world)
        self.app = app
        self.params = list(params if params is not None else ecs_apps.DEFAULT_PARAMS[app])
        p, pp, pn = _ids(self.params)
        _check(self.lib.mdr_app_setup(self.h, app, pp, pn), "mdr_app_setup")
        self._describe()
        if init:
            for (arch, col), data in ecs_apps.initial_columns(app, num_worlds, self.params, first_world).items():
                self.write_column_all(arch, col, data)

    # whole-state accessor used by the parity tests
    def read_world_columns(self):
        """{(arch, col): [W, stride, bytes]} of every column, one device copy each."""
        return {(a, c): self.read_column_all(a, c) for a in range(self.num_archetypes)
                for c in range(len(self.col_bytes[a]))}


def static_map_build(keys, values, lib=None):
    """StaticIntegerMap<128>(keys, values) of the reference: (key_storage[128], value_storage[128]).
    Host code of the library (csrc/static_map.hpp); raises where the reference FATALs."""
    lib = lib or _lib()
    k = np.ascontiguousarray(keys, np.uint32)
    v = np.ascontiguousarray(values, np.uint32)
    ks, vs = np.zeros(128, np.uint32), np.zeros(128, np.uint32)
    _check(lib.mdr_static_map_build(k.ctypes.data, v.ctypes.data, len(k), ks.ctypes.data, vs.ctypes.data),
           "mdr_static_map_build")
    return ks, vs


def static_map_lookup(key_storage, value_storage, key, lib=None):
    lib = lib or _lib()
    return int(lib.mdr_static_map_lookup(key_storage.ctypes.data, value_storage.ctypes.data, int(key)))
