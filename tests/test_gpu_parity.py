"""Parity of the CUDA world step (through the C ABI) against the oracle, the
golden fixtures of the compiled reference, and size-independent properties at
BASELINE.json's full size. Bit-exact: positions, rotations, velocities, pair
lists, contact manifolds."""
import numpy as np
import pytest

from gpu_ecs_madrona_b200 import phys, scenes
from tests.golden.make_golden import CASES
from tests.util import assert_state_equal, bits, contacts_equal, load_golden, worlds_equal

pytestmark = pytest.mark.gpu

SCENES = {
    "collisions64": lambda: scenes.collisions64(64),
    "collisions16": lambda: scenes.collisions64(33, num_cubes=16),
    "tumble24": lambda: scenes.tumble(48),
    "tumble40": lambda: scenes.tumble(16, num_bodies=40, spread=1.0, seed=3),
    "tumble12_noplane": lambda: scenes.tumble(16, num_bodies=12, seed=7, spread=0.8, with_plane=False),
    "jointed": lambda: scenes.jointed(37),
}


@pytest.mark.parametrize("threads", [32, 64, 128])
@pytest.mark.parametrize("name", sorted(SCENES))
def test_cuda_matches_oracle(cuda_lib, port_oracle, name, threads):
    sc = SCENES[name]()
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, block_threads=threads)
    for st in range(40):
        o.step(1, 8)
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"{name} T={threads} step {st}")
    for col in (phys.COL_SUBSTEP_PREV, phys.COL_PRESOLVE_POS, phys.COL_PRESOLVE_VEL):
        assert np.array_equal(e.read(col), o.read(col)), col


@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_matches_reference_golden(cuda_lib, port_oracle, name):
    fn, kw, snaps, pw = CASES[name]
    g = load_golden(name)
    sc = fn(**kw)
    e = phys.PhysExecutor(sc)
    step = 0
    for snap in snaps:
        e.run(snap - step)
        step = snap
        # every world: the fixtures are the reference build whose manifold slots are defined (make_golden.py)
        assert_state_equal(e.state(), (g[f"pos_{snap}"], g[f"rot_{snap}"], g[f"vel_{snap}"]),
                           None, f"{name} step {snap}")
    e2 = phys.PhysExecutor(sc, smem_contacts=8)    # exercises the global overflow slots
    pr = e2.probe_step(pw)
    assert np.array_equal(pr["pairs"], g["probe_pairs"])
    assert np.array_equal(pr["num_contacts"], g["probe_num_contacts"])
    slot, leaf, order = e2.read_bvh()
    assert np.array_equal(bits(slot[pw]), bits(g["probe_slot_boxes"]))
    assert np.array_equal(bits(leaf[pw]), bits(g["probe_leaf_aabbs"]))
    rank = np.empty(sc.bodies_per_world, np.int32)
    rank[order[pw].astype(np.int64)] = np.arange(sc.bodies_per_world)
    assert np.array_equal(rank, g["probe_ref_rank"])
    for s, c in enumerate(pr["contacts"]):
        assert contacts_equal(c, g[f"probe_contacts_{s}"]), f"substep {s}"


def test_cuda_hinges_and_joint_table_updates(cuda_lib, port_oracle):
    """Hinge joints while the reference keeps the state finite (scenes.jointed), then the joint
    table replaced and removed through mdr_phys_set_joints between steps."""
    sc = scenes.jointed(21, hinges=True)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc)
    for st in range(8):
        o.step()
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"hinges step {st}")
    plain = scenes.jointed(21)
    e2 = phys.PhysExecutor(sc)
    e2.set_joints(plain.joints, plain.joint_counts)
    o2 = port_oracle.Oracle(plain, "port")
    e2.run(30)
    o2.step(30)
    assert_state_equal(e2.state(), o2.state(), None, "joint table replaced")
    e2.set_joints(None)
    bare = scenes.jointed(21)
    bare.joints = bare.joint_counts = None
    o3 = port_oracle.Oracle(bare, "port")
    for c in (phys.COL_POSITION, phys.COL_ROTATION, phys.COL_VELOCITY):
        o3.write(c, e2.read(c))
    o3.flag_rebuild(); e2.flag_rebuild()
    e2.run(5)
    o3.step(5)
    assert_state_equal(e2.state(), o3.state(), None, "joints removed")
    bad = plain.joints.copy()
    bad["row1"][3, 0] = -1
    with pytest.raises(phys.MadronaError):
        e2.set_joints(bad, plain.joint_counts)


def test_cuda_drift_100_steps(cuda_lib, port_oracle):
    """north_star: within a stated drift bound after 100 steps. The bound here is 0."""
    sc = scenes.collisions64(32)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc)
    o.step(100, 8)
    e.run(100)
    assert_state_equal(e.state(), o.state(), None, "100 steps")


@pytest.mark.parametrize("name", ["tumble48", "tumble40", "hull_pile"])
def test_cuda_matches_defined_reference_100_steps(cuda_lib, ref_oracle, name):
    """Hull-hull scenes, 100 steps, EVERY world against the compiled reference itself: the build whose locals are
    defined (oracle/_ref/libmadrona_ref_defined.so), so the worlds in which buildFaceContactManifold returns a
    slot it never wrote (11 of 48 / 16 / 32 here) are compared too, not masked."""
    if not ref_oracle.available("reference_defined"):
        pytest.skip("oracle/_ref/libmadrona_ref_defined.so did not travel")
    sc = {"tumble48": lambda: scenes.tumble(48),
          "tumble40": lambda: scenes.tumble(16, num_bodies=40, spread=1.0, seed=3),
          "hull_pile": lambda: scenes.hull_pile(32)}[name]()
    d = ref_oracle.Oracle(sc, "reference_defined")
    e = phys.PhysExecutor(sc)
    for st in range(0, 100, 10):
        d.step(10, 8)
        e.run(10)
        assert_state_equal(e.state(), d.state(), None, f"{name} step {st + 10}")


@pytest.mark.parametrize("name", ["collisions64", "tumble24", "jointed", "few_worlds"])
def test_cuda_batched_steps_equal_single_steps(cuda_lib, monkeypatch, name):
    """run(n > 1) is ONE launch that walks (step, world group) tasks, a group's step waiting for its previous one
    wherever that ran (phys_kernel.cuh). Against n launches of one step (MDR_PHYS_NO_BATCH=1): every column."""
    make = {"collisions64": lambda: scenes.collisions64(1500), "tumble24": lambda: scenes.tumble(700),
            "jointed": lambda: scenes.jointed(333), "few_worlds": lambda: scenes.collisions64(3)}[name]
    batch = phys.PhysExecutor(make())
    monkeypatch.setenv("MDR_PHYS_NO_BATCH", "1")
    single = phys.PhysExecutor(make())
    monkeypatch.delenv("MDR_PHYS_NO_BATCH")
    for steps in (7, 1, 40):
        batch.run(steps)
        single.run(steps)
        assert batch.last_launches() == 1 and single.last_launches() == steps
        for col in (phys.COL_POSITION, phys.COL_ROTATION, phys.COL_VELOCITY, phys.COL_SUBSTEP_PREV,
                    phys.COL_PRESOLVE_POS, phys.COL_PRESOLVE_VEL):
            assert np.array_equal(batch.read(col), single.read(col)), (steps, col)
        for a, b in zip(batch.read_bvh(), single.read_bvh()):
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), steps
        for a, b in zip(batch.stats(), single.stats()):
            assert np.array_equal(a, b), steps
    batch.close()
    single.close()


def test_cuda_ragged_and_empty_worlds(cuda_lib, port_oracle):
    full = scenes.tumble(6, num_bodies=15)
    rows = np.array([16, 9, 1, 0, 5, 16], np.int32)
    full.rows_per_world = rows
    e = phys.PhysExecutor(full)
    e.run(12)
    got = e.state()
    for w, n in enumerate(rows):
        if n == 0:
            continue
        sub = scenes.tumble(6, num_bodies=15)
        sl = slice(w, w + 1)
        one = scenes.Scene(1, int(n), sub.num_substeps, sub.max_contacts, sub.delta_t, sub.gravity,
                           sub.objects, sub.pos[sl, :n], sub.rot[sl, :n], sub.scale[sl, :n],
                           sub.vel[sl, :n], sub.obj_ids[sl, :n], sub.response[sl, :n],
                           sub.ext_force[sl, :n], sub.ext_torque[sl, :n])
        o = port_oracle.Oracle(one, "port")
        o.step(12)
        for a, b in zip(got, o.state()):
            assert np.array_equal(bits(a[w, :n]), bits(b[0])), f"world {w}"
    assert np.array_equal(got[0][3], np.ascontiguousarray(full.pos[3], np.float32))


def test_cuda_pair_overflow_reports(cuda_lib):
    sc = scenes.tumble(2, num_bodies=40, spread=0.6, seed=5, max_contacts=64)
    e = phys.PhysExecutor(sc)
    with pytest.raises(phys.MadronaError) as ei:
        e.run()
    assert ei.value.code == phys.MDR_EOVERFLOW


def test_cuda_rebuild_flag(cuda_lib, port_oracle):
    sc = scenes.tumble(8, num_bodies=20)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc)
    o.step(10); e.run(10)
    o.flag_rebuild(); e.flag_rebuild()
    for st in range(10):
        o.step(); e.run()
        assert_state_equal(e.state(), o.state(), None, f"after rebuild step {st}")


def test_full_size_properties(cuda_lib, port_oracle):
    """collisions-64 at 16 384 worlds (BASELINE.json configs[2]): (1) a sample
    of worlds spread over the whole grid equals the oracle bit for bit, (2) the
    result does not depend on the block size / which block stepped a world,
    (3) sharding by world_offset reproduces the same worlds, (4) state stays
    finite and cubes stay above the plane."""
    W, steps = 16384, 20
    sc = scenes.collisions64(W)
    e = phys.PhysExecutor(sc)
    e.run(steps)
    got = e.state()
    idx = np.unique(np.concatenate([np.arange(0, W, 257), [W - 1]]))
    for w in idx[:: max(1, len(idx) // 24)]:
        one = scenes.collisions64(1, world_offset=int(w))
        o = port_oracle.Oracle(one, "port")
        o.step(steps)
        for a, b in zip(got, o.state()):
            assert np.array_equal(bits(a[w]), bits(b[0])), f"world {w}"
    e128 = phys.PhysExecutor(sc, block_threads=128)
    e128.run(steps)
    assert worlds_equal(got, e128.state()).all()
    shard = scenes.collisions64(1024, world_offset=8192)
    es = phys.PhysExecutor(shard)
    es.run(steps)
    for a, b in zip(got, es.state()):
        assert np.array_equal(bits(a[8192:8192 + 1024]), bits(b))
    assert np.isfinite(got[0]).all() and np.isfinite(got[2]).all()
    assert (got[0][:, 1:, 2] > 0.4).all()
    pairs, contacts = e.stats()
    assert pairs.min() >= 1 and contacts.max() <= sc.max_contacts


def test_exported_column_pointer_is_in_place(cuda_lib):
    """getExported semantics: the device pointer aliases the table (zero copy)."""
    import torch
    sc = scenes.collisions64(8)
    e = phys.PhysExecutor(sc)
    e.run(2)
    ptr = e.column_ptr(phys.COL_POSITION)
    assert ptr
    stride = (sc.bodies_per_world + 3) // 4 * 4
    n = sc.num_worlds * stride * 3
    buf = (torch.empty(0, dtype=torch.float32, device="cuda:0"))
    import ctypes
    import glob
    import os
    host = np.empty(n, np.float32)
    cands = glob.glob("/usr/local/cuda/lib64/libcudart.so*") + glob.glob(
        os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
    cudart = ctypes.CDLL(sorted(cands)[0])
    rc = cudart.cudaMemcpy(ctypes.c_void_p(host.ctypes.data), ctypes.c_void_p(ptr),
                           ctypes.c_size_t(4 * n), ctypes.c_int(2))
    assert rc == 0
    pos = host.reshape(sc.num_worlds, stride, 3)[:, :sc.bodies_per_world]
    assert np.array_equal(bits(pos), bits(e.state()[0]))
    del buf


@pytest.mark.parametrize("chunks", [1, 2, 5])
def test_step_io_pipeline_matches_plain_step(cuda_lib, chunks):
    """mdr_phys_step_io_async (copy-in, step and copy-out pipelined over world ranges)
    gives the bits of write -> step -> read, forces included."""
    sc = scenes.collisions64(3000, num_cubes=7)      # several resident waves of worlds
    a, b = phys.PhysExecutor(sc), phys.PhysExecutor(sc)
    rng = np.random.default_rng(5)
    in_cols = (phys.COL_EXT_FORCE, phys.COL_EXT_TORQUE)
    out_cols = (phys.COL_POSITION, phys.COL_ROTATION, phys.COL_VELOCITY)
    ins = [(c, b.pinned(c)) for c in in_cols]
    outs = [(c, b.pinned(c)) for c in out_cols]
    for st in range(6):
        for c, arr in ins:
            f = rng.uniform(-3, 3, (sc.num_worlds, b.row_stride(), 3)).astype(np.float32)
            arr[:] = f.view(np.uint8).reshape(arr.shape)
            a.write_raw_async(c, arr)
        a.run(1)
        b.step_io_async(ins, outs, chunks)
        b.sync()
        for c, arr in outs:
            want = a.read(c)
            got = arr.view(np.float32).reshape(sc.num_worlds, b.row_stride(), -1)[:, :sc.bodies_per_world]
            assert np.array_equal(bits(got), bits(want)), (st, c)
    assert_state_equal(b.state(), a.state(), None, "after pipelined steps")


def test_partial_reset_rebuilds_only_the_named_worlds(cuda_lib, port_oracle):
    """An RL episode reset of a subset of worlds (RigidBodyPhysicsSystem::reset is per world,
    physics.cpp:1038-1043): their poses go back to the start and only THEIR BVHs are rebuilt; every
    other world keeps its grown slot boxes and traversal order, hence its pair order and its bits."""
    sc = scenes.collisions64(40)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc)
    o.step(25, 8)
    e.run(25)
    assert_state_equal(e.state(), o.state(), None, "before the reset")
    subset = [1, 5, 6, 7, 20, 39]
    vel0 = np.ascontiguousarray(sc.vel, np.float32)
    for col, init in ((phys.COL_POSITION, sc.pos), (phys.COL_ROTATION, sc.rot), (phys.COL_VELOCITY, vel0)):
        full = o.read_f32(col).copy()
        for w in subset:
            full[w] = np.asarray(init, np.float32)[w].reshape(full[w].shape)
            e.write_world(col, w, np.asarray(init, np.float32)[w])
        o.write(col, full)
    o.flag_rebuild(subset)
    e.reset_worlds(subset)
    slot_before = e.read_bvh()[0]
    for st in range(30):
        o.step(1, 8)
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"step {st} after the reset")
    # a world that was not reset still carries boxes at least as large as before (grow-only)
    slot_after = e.read_bvh()[0]
    others = [w for w in range(40) if w not in subset]
    assert (slot_after[others][..., 3:] >= slot_before[others][..., 3:]).all()
    assert (slot_after[others][..., :3] <= slot_before[others][..., :3]).all()
    assert e.lib.mdr_phys_reset_worlds(e.h, np.array([40], np.int32).ctypes.data, 1) == phys.MDR_EINVAL


def test_step_io_with_a_flagged_rebuild_reads_the_new_poses(cuda_lib, port_oracle):
    """mdr_phys_step_io_async with Position / Rotation among the INPUT columns on a step whose BVH rebuild is
    flagged (an episode reset through the pipelined host I/O call): the rebuild must see the poses that come
    in with the same call."""
    sc = scenes.collisions64(300)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc)
    o.step(12, 8)
    e.run(12)
    rng = np.random.default_rng(3)
    pos = o.read_f32(phys.COL_POSITION).copy()
    pos[:, 1:, :2] += rng.uniform(-0.05, 0.05, pos[:, 1:, :2].shape).astype(np.float32)
    pos[:, 1:, 2] += np.float32(0.75)
    o.write(phys.COL_POSITION, pos)
    o.flag_rebuild()
    o.step(1, 8)
    pin = e.pinned(phys.COL_POSITION)
    pin[:] = 0
    pin.view(np.float32).reshape(300, e.row_stride(), 3)[:, :sc.bodies_per_world] = pos
    pout = e.pinned(phys.COL_VELOCITY)
    e.flag_rebuild()
    e.step_io_async([(phys.COL_POSITION, pin)], [(phys.COL_VELOCITY, pout)], 4)
    e.sync()
    assert_state_equal(e.state(), o.state(), None, "step with rebuild + pipelined inputs")
    o.step(5, 8)
    e.run(5)
    assert_state_equal(e.state(), o.state(), None, "after")
    e.close(); o.close()


def test_object_id_outside_the_table_is_an_error_code(cuda_lib):
    """an ObjectID the caller writes through a raw / tensor path used to be dereferenced on the device.
    It is flagged, treated as object 0 for the step, and mdr_phys_sync answers MDR_EINVAL; the executor stays usable."""
    sc = scenes.collisions64(8)
    e = phys.PhysExecutor(sc)
    e.run(2)
    ids = e.pinned(phys.COL_OBJECT_ID)
    good = np.zeros((8, e.row_stride()), np.int32)
    good[:, :sc.bodies_per_world] = sc.obj_ids
    bad = good.copy()
    bad[3, 5] = 77
    ids.view(np.int32).reshape(8, e.row_stride())[:] = bad
    e.write_raw_async(phys.COL_OBJECT_ID, ids)
    with pytest.raises(phys.MadronaError) as ei:
        e.run(1)
    assert ei.value.code == phys.MDR_EINVAL
    ids.view(np.int32).reshape(8, e.row_stride())[:] = good
    e.write_raw_async(phys.COL_OBJECT_ID, ids)
    e.run(3)
    assert np.isfinite(e.state()[0]).all()
    with pytest.raises(phys.MadronaError):          # the checked path refuses it up front
        e.write(phys.COL_OBJECT_ID, np.full((8, sc.bodies_per_world), 9, np.int32))
    e.close()
