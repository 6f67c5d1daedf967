"""TEST INFRASTRUCTURE ONLY: ctypes loader for the ECS / task-graph checkers
behind oracle/ecs_api.h (`_ref/libmadrona_ecs_ref.so` = the compiled reference,
kind "reference"; `libmadrona_ecs_port.so` = our scalar restatement, kind "port").
Imported only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libmadrona_ecs_ref.so")
PORT_SO = os.path.join(HERE, "libmadrona_ecs_port.so")

APP_SCRIPT, APP_TASKGRAPH, APP_FANOUT, APP_FVS = range(4)
OP_CREATE, OP_DESTROY, OP_CLEAR, OP_TMP, OP_CLEAR_TMP = 1, 2, 3, 4, 5
OP_DTYPE = np.dtype([("op", "<i4"), ("a", "<i4"), ("b", "<i4")])

_libs = {}


def available(kind):
    return os.path.exists(REF_SO if kind == "reference" else PORT_SO)


def _load(kind):
    if kind in _libs:
        return _libs[kind]
    path = REF_SO if kind == "reference" else PORT_SO
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} is not built (make -C oracle)")
    lib = C.CDLL(path)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.eorc_kind.restype = C.c_char_p
    lib.eorc_create.restype = vp
    lib.eorc_create.argtypes = [i32, i32, i32, vp, i32]
    lib.eorc_destroy.argtypes = [vp]
    lib.eorc_step.argtypes = [vp, i32, i32]
    lib.eorc_script.argtypes = [vp, i32, vp, i32]
    lib.eorc_script.restype = i32
    for name, args in (("eorc_num_archetypes", [vp]), ("eorc_num_cols", [vp, i32]),
                       ("eorc_col_bytes", [vp, i32, i32]), ("eorc_num_rows", [vp, i32, i32]),
                       ("eorc_id_count", [vp, i32])):
        getattr(lib, name).argtypes = args
        getattr(lib, name).restype = i32
    lib.eorc_read_column.argtypes = [vp, i32, i32, i32, vp]
    lib.eorc_write_column.argtypes = [vp, i32, i32, i32, vp]
    lib.eorc_read_idmap.argtypes = [vp, i32, vp, vp, vp, vp]
    lib.eorc_export.argtypes = [vp, i32, vp, i64]
    lib.eorc_export.restype = i64
    _libs[kind] = lib
    return lib


class EcsOracle:
    def __init__(self, kind, app, num_worlds, params=(), first_world=0):
        self.lib = _load(kind)
        self.kind, self.app, self.num_worlds = kind, app, num_worlds
        p = np.asarray(list(params), np.int32)
        self.h = self.lib.eorc_create(app, num_worlds, first_world, p.ctypes.data if len(p) else None, len(p))
        if not self.h:
            raise RuntimeError("eorc_create failed")
        self.num_archetypes = self.lib.eorc_num_archetypes(self.h)
        self.col_bytes = [[self.lib.eorc_col_bytes(self.h, a, c) for c in range(self.lib.eorc_num_cols(self.h, a))]
                          for a in range(self.num_archetypes)]

    def close(self):
        if self.h:
            self.lib.eorc_destroy(self.h)
            self.h = None

    def step(self, n=1, threads=1):
        self.lib.eorc_step(self.h, n, threads)

    def script(self, world, ops):
        ops = np.ascontiguousarray(ops, OP_DTYPE)
        rc = self.lib.eorc_script(self.h, world, ops.ctypes.data, len(ops))
        if rc != 0:
            raise RuntimeError("eorc_script failed")

    def num_rows(self, world, arch):
        return self.lib.eorc_num_rows(self.h, world, arch)

    def row_counts(self):
        return np.array([[self.num_rows(w, a) for a in range(self.num_archetypes)]
                         for w in range(self.num_worlds)], np.int32)

    def read_column(self, world, arch, col):
        n, nb = self.num_rows(world, arch), self.col_bytes[arch][col]
        out = np.zeros((n, nb), np.uint8)
        if n:
            self.lib.eorc_read_column(self.h, world, arch, col, out.ctypes.data)
        return out

    def write_column(self, world, arch, col, data):
        n, nb = self.num_rows(world, arch), self.col_bytes[arch][col]
        d = np.ascontiguousarray(data).view(np.uint8).reshape(n, nb)
        if n:
            self.lib.eorc_write_column(self.h, world, arch, col, d.ctypes.data)

    def idmap(self, world):
        n = self.lib.eorc_id_count(self.h, world)
        a, b = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32)
        gen, state = np.zeros(max(n, 1), np.uint32), np.zeros(10, np.int32)
        self.lib.eorc_read_idmap(self.h, world, a.ctypes.data, b.ctypes.data, gen.ctypes.data, state.ctypes.data)
        return a[:n], b[:n], gen[:n], state

    def export(self, slot, max_bytes=1 << 26):
        out = np.zeros(max_bytes, np.uint8)
        n = self.lib.eorc_export(self.h, slot, out.ctypes.data, max_bytes)
        if n < 0:
            raise RuntimeError("eorc_export failed")
        return out[:n].copy()


def make_app_oracle(kind, app, num_worlds, params=None, first_world=0):
    """A checker for one of the task-graph worlds with the SAME initial columns the CUDA executor gets
    (gpu_ecs_madrona_b200.ecs_apps.initial_columns: host-side input data only)."""
    from gpu_ecs_madrona_b200 import ecs_apps
    params = list(params if params is not None else ecs_apps.DEFAULT_PARAMS[app])
    o = EcsOracle(kind, app, num_worlds, params, first_world)
    for (arch, col), data in ecs_apps.initial_columns(app, num_worlds, params, first_world).items():
        for w in range(num_worlds):
            o.write_column(w, arch, col, data[w])
    return o
