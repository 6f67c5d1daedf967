"""Host mirror of the reference's SortArchetypeNode for tables stored as one global SoA
(src/mw/device/sort_archetype.cpp:1267-1392; include/madrona/taskgraph.hpp SortArchetypeNode):
`TableSort(column_bytes, key_column, num_worlds).run(columns_in, columns_out, num_rows)` over the
mdr_sort_* C ABI (include/madrona_b200.h). Columns are device buffers handed over as torch
tensors (any dtype, contiguous, 16-byte aligned) or raw device pointers; torch only owns memory.
There is no CPU path: the CUDA library must be built."""
import ctypes as C

import numpy as np

from . import phys


def _bind(lib):
    vp, i32 = C.c_void_p, C.c_int32
    lib.mdr_sort_num_passes.argtypes = [i32]
    lib.mdr_sort_create.argtypes = [i32, i32, i32, vp, i32, i32, C.POINTER(vp)]
    lib.mdr_sort_destroy.argtypes = [vp]
    lib.mdr_sort_destroy.restype = None
    lib.mdr_sort_run.argtypes = [vp, vp, vp, i32, vp, i32, vp]
    lib.mdr_sort_last_ms.argtypes = [vp, vp]
    lib.mdr_sort_last_ms.restype = C.c_float
    return lib


def _ptr(x):
    return int(x) if isinstance(x, int) else int(x.data_ptr())


class TableSort:
    """Sorts every column of a shared table by a u32 key column (column 0 = Entity {gen, id}).
    ``num_worlds > 0``: the key is the WorldID column: rows with WorldID -1 are dropped and
    ``run`` returns the new row count; ``num_worlds = 0``: a generic 32-bit key, all rows kept."""

    def __init__(self, max_rows, column_bytes, key_column, num_worlds, gpu_id=0, lib=None):
        self.lib = _bind(lib if lib is not None else phys.load_library())
        self.column_bytes = [int(b) for b in column_bytes]
        cb = (C.c_int32 * len(self.column_bytes))(*self.column_bytes)
        h = C.c_void_p()
        rc = self.lib.mdr_sort_create(gpu_id, max_rows, len(self.column_bytes), cb, key_column, num_worlds,
                                      C.byref(h))
        if rc != phys.MDR_OK:
            raise phys.MadronaError(rc, "mdr_sort_create")
        self.h = h
        self.passes = self.lib.mdr_sort_num_passes(num_worlds)

    def run(self, columns_in, columns_out, num_rows, entity_row=None):
        n = len(self.column_bytes)
        assert len(columns_in) == n and len(columns_out) == n
        pin = (C.c_void_p * n)(*[_ptr(c) for c in columns_in])
        pout = (C.c_void_p * n)(*[_ptr(c) for c in columns_out])
        new_rows = C.c_int32(0)
        er, ne = (None, 0) if entity_row is None else (_ptr(entity_row), int(entity_row.numel()))
        rc = self.lib.mdr_sort_run(self.h, pin, pout, num_rows, er, ne, C.byref(new_rows))
        if rc != phys.MDR_OK:
            raise phys.MadronaError(rc, "mdr_sort_run")
        return int(new_rows.value)

    def last_ms(self):
        n = C.c_int32(0)
        ms = self.lib.mdr_sort_last_ms(self.h, C.byref(n))
        return float(ms), int(n.value)

    def close(self):
        if getattr(self, "h", None):
            self.lib.mdr_sort_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def churned_table(num_worlds, rows_per_world, column_bytes, dead_fraction=0.1, shuffle_fraction=0.2, seed=0):
    """A synthetic shared table in the state the reference sorts it in: rows appended in creation
    order (mostly grouped by world), a fraction created late (out of world order), a fraction
    destroyed (WorldID -1). Returns numpy columns: [0] Entity {gen, id}, [1] WorldID, then
    payload columns of the given byte sizes filled with a function of the row's entity id."""
    rng = np.random.default_rng(seed)
    n = num_worlds * rows_per_world
    world = np.repeat(np.arange(num_worlds, dtype=np.int32), rows_per_world)
    late = rng.random(n) < shuffle_fraction
    order = np.concatenate([np.nonzero(~late)[0], rng.permutation(np.nonzero(late)[0])])
    world = world[order]
    world[rng.random(n) < dead_fraction] = -1
    ids = rng.permutation(n).astype(np.int32)
    ent = np.stack([rng.integers(0, 4, n, dtype=np.int32), ids], axis=1)
    cols = [ent, world]
    for c, b in enumerate(column_bytes[2:]):
        w = b // 4
        cols.append((ids[:, None].astype(np.uint32) * np.uint32(2654435761)
                     + np.arange(w, dtype=np.uint32)[None] * np.uint32(40503) + np.uint32(c)).astype(np.uint32))
    return cols
