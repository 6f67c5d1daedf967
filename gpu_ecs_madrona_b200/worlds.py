"""Host-side handles of the OUT-OF-TREE example worlds (examples/*.cu): user systems written against
include/madrona/b200 (the reference's registerTypes / setupTasks / ParallelForNode surface), compiled
by nvcc into their own shared objects and linked to libmadrona_b200.so. The tests drive them like the
reference's example mains drive theirs (examples/simple_taskgraph/mw_cpu.cpp:29-42): create, run, read
exported columns. State is inspected through the library's C ABI on the world's own mdr_ecs handle."""
import ctypes as C
import os

import numpy as np

from . import phys
from .ecs import EcsRegistry

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(rel):
    path = os.path.join(ROOT, "examples", rel)
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: build it with `bash examples/build.sh` (nvcc, sm_100a)")
    phys.load_library()          # the world links against it: make sure it is this tree's build
    return C.CDLL(path)


class SimpleTaskgraphWorld:
    """examples/simple_taskgraph/simple.cu: positions [W, num_spheres + 1, 3] f32 (spheres, then the agent)."""

    def __init__(self, positions, first_world=0):
        lib = _load("simple_taskgraph/libsimple_world.so")
        lib.simple_create.restype = C.c_void_p
        lib.simple_create.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
        lib.simple_run.argtypes = [C.c_void_p, C.c_int32]
        lib.simple_ecs.restype = C.c_void_p
        lib.simple_ecs.argtypes = [C.c_void_p]
        lib.simple_destroy.argtypes = [C.c_void_p]
        self.lib = lib
        pos = np.ascontiguousarray(positions, np.float32)
        W, n1, _ = pos.shape
        self.h = lib.simple_create(W, n1 - 1, first_world, pos.ctypes.data)
        self.ecs = EcsRegistry.attach(lib.simple_ecs(self.h), W)

    def run(self, steps=1):
        self.lib.simple_run(self.h, steps)

    def close(self):
        if self.h:
            self.ecs.close()
            self.lib.simple_destroy(self.h)
            self.h = None


class ChurnWorld:
    """examples/churn/churn.cu: immediate-mode makeEntityNow / destroyEntityNow churn."""

    def __init__(self, num_worlds, num_spawners=24, max_bullets=640, first_world=0):
        lib = _load("churn/libchurn_world.so")
        lib.churn_create.restype = C.c_void_p
        lib.churn_create.argtypes = [C.c_int32] * 4
        lib.churn_run.argtypes = [C.c_void_p, C.c_int32]
        lib.churn_ecs.restype = C.c_void_p
        lib.churn_ecs.argtypes = [C.c_void_p]
        lib.churn_destroy.argtypes = [C.c_void_p]
        self.lib = lib
        self.h = lib.churn_create(num_worlds, num_spawners, max_bullets, first_world)
        self.ecs = EcsRegistry.attach(lib.churn_ecs(self.h), num_worlds)

    def run(self, steps=1):
        self.lib.churn_run(self.h, steps)

    def close(self):
        if self.h:
            self.ecs.close()
            self.lib.churn_destroy(self.h)
            self.h = None


class ForceFieldWorld:
    """examples/force_field/force_field.cu: a scenes.Scene stepped as forceSystem -> physics -> obsSystem."""
    NUM_PHYS_COLS = 13

    def __init__(self, scene, stiffness, gpu_id=0):
        lib = _load("force_field/libforce_field_world.so")
        lib.ff_create.restype = C.c_void_p
        lib.ff_create.argtypes = [C.c_int32, C.c_int32, C.c_float, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
        lib.ff_run.argtypes = [C.c_void_p, C.c_int32]
        lib.ff_ecs.restype = C.c_void_p
        lib.ff_ecs.argtypes = [C.c_void_p]
        lib.ff_destroy.argtypes = [C.c_void_p]
        self.lib = lib
        cfg = phys.MdrPhysConfig(scene.num_worlds, scene.bodies_per_world, scene.num_substeps, scene.max_contacts,
                                 scene.delta_t, (C.c_float * 3)(*scene.gravity), gpu_id, 0, 0, 0)
        objs, verts, words = phys.pack_objects(scene.objects)
        self.h = lib.ff_create(scene.num_worlds, scene.bodies_per_world, stiffness, C.byref(cfg), objs,
                               len(scene.objects), verts.ctypes.data, words.ctypes.data)
        self.ecs = EcsRegistry.attach(lib.ff_ecs(self.h), scene.num_worlds, gpu_id)
        vel = np.ascontiguousarray(scene.vel, np.float32)
        for col, arr, dt in ((phys.COL_POSITION, scene.pos, np.float32), (phys.COL_ROTATION, scene.rot, np.float32),
                             (phys.COL_SCALE, scene.scale, np.float32), (phys.COL_VELOCITY, vel, np.float32),
                             (phys.COL_OBJECT_ID, scene.obj_ids, np.int32), (phys.COL_RESPONSE_TYPE, scene.response, np.uint32),
                             (phys.COL_EXT_FORCE, scene.ext_force, np.float32), (phys.COL_EXT_TORQUE, scene.ext_torque, np.float32)):
            self.ecs.write_column_all(0, col, np.ascontiguousarray(arr, dt))

    def run(self, steps=1):
        self.lib.ff_run(self.h, steps)

    def column(self, col, dtype=np.float32):
        n = self.ecs.row_counts()[:, 0].max()
        return self.ecs.read_column_all(0, col, int(n)).view(dtype)

    def close(self):
        if self.h:
            self.ecs.close()
            self.lib.ff_destroy(self.h)
            self.h = None
