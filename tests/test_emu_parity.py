"""CPU check of the step's RESTRUCTURED algorithm: the device source compiled
for the host (tests/emu, a test tool) must equal the oracle bit for bit. This
is what `-m "not gpu"` can say about the kernels; tests/test_gpu_parity.py runs
the same comparisons through the CUDA library on the B200."""
import numpy as np
import pytest

from gpu_ecs_madrona_b200 import phys, scenes
from tests.golden.make_golden import CASES
from tests.util import assert_state_equal, bits, contacts_equal, load_golden

SCENES = {
    "collisions64": lambda: scenes.collisions64(3),
    "collisions16": lambda: scenes.collisions64(5, num_cubes=16),
    "tumble24": lambda: scenes.tumble(8),
    "tumble40": lambda: scenes.tumble(4, num_bodies=40, spread=1.0, seed=3),
    "tumble12_noplane": lambda: scenes.tumble(4, num_bodies=12, seed=7, spread=0.8, with_plane=False),
    "jointed": lambda: scenes.jointed(6),
}


@pytest.mark.parametrize("name", sorted(SCENES))
def test_emu_matches_oracle(port_oracle, emu_lib, name):
    sc = SCENES[name]()
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    for st in range(30):
        o.step()
        e.run()
        assert_state_equal(e.state(), o.state(), None, f"{name} step {st}")
    for col in (phys.COL_SUBSTEP_PREV, phys.COL_PRESOLVE_POS, phys.COL_PRESOLVE_VEL):
        assert np.array_equal(e.read(col), o.read(col)), col


def test_emu_hinges_match_oracle(port_oracle, emu_lib):
    """Hinge joints (physics.cpp:511-535,608-618), while the reference keeps the state finite
    (see scenes.jointed); then joints replaced / removed through mdr_phys_set_joints."""
    sc = scenes.jointed(5, hinges=True)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    for st in range(8):
        o.step()
        e.run()
        assert all(np.isfinite(a).all() for a in o.state())
        assert_state_equal(e.state(), o.state(), None, f"hinges step {st}")
    plain = scenes.jointed(5)
    e2 = phys.PhysExecutor(sc, lib=emu_lib)
    e2.set_joints(plain.joints, plain.joint_counts)     # swap the hinge table for the fixed one
    o2 = port_oracle.Oracle(plain, "port")
    e2.run(10)
    o2.step(10)
    assert_state_equal(e2.state(), o2.state(), None, "joint table replaced")
    e3 = phys.PhysExecutor(plain, lib=emu_lib)
    e3.set_joints(None)
    plain.joints = plain.joint_counts = None
    o3 = port_oracle.Oracle(plain, "port")
    e3.run(10)
    o3.step(10)
    assert_state_equal(e3.state(), o3.state(), None, "joints removed")


def test_emu_set_joints_rejects_bad_arguments(emu_lib):
    sc = scenes.jointed(2)
    e = phys.PhysExecutor(sc, lib=emu_lib)
    bad = sc.joints.copy()
    bad["row2"][1, 0] = sc.bodies_per_world
    with pytest.raises(phys.MadronaError):
        e.set_joints(bad, sc.joint_counts)
    bad = sc.joints.copy()
    bad["type"][0, 2] = 2
    with pytest.raises(phys.MadronaError):
        e.set_joints(bad, sc.joint_counts)
    with pytest.raises(phys.MadronaError):
        e.set_joints(sc.joints, np.full(2, sc.joints.shape[1] + 1, np.int32))


@pytest.mark.parametrize("name", sorted(CASES))
def test_emu_probe_matches_golden(emu_lib, port_oracle, name):
    fn, kw, snaps, pw = CASES[name]
    g = load_golden(name)
    sc = fn(**kw)
    e = phys.PhysExecutor(sc, lib=emu_lib)
    pr = e.probe_step(pw)
    assert np.array_equal(pr["pairs"], g["probe_pairs"])
    assert np.array_equal(pr["num_contacts"], g["probe_num_contacts"])
    slot, leaf, order = e.read_bvh()
    assert np.array_equal(bits(slot[pw]), bits(g["probe_slot_boxes"]))
    assert np.array_equal(bits(leaf[pw]), bits(g["probe_leaf_aabbs"]))
    rank = np.empty(sc.bodies_per_world, np.int32)
    rank[order[pw].astype(np.int64)] = np.arange(sc.bodies_per_world)
    assert np.array_equal(rank, g["probe_ref_rank"])
    for s, c in enumerate(pr["contacts"]):
        assert contacts_equal(c, g[f"probe_contacts_{s}"]), f"substep {s}"


def test_emu_ragged_and_empty_worlds(port_oracle, emu_lib):
    """Worlds with fewer rows than the capacity, one with a single body, one empty."""
    full = scenes.tumble(6, num_bodies=15)
    rows = np.array([16, 9, 1, 0, 5, 16], np.int32)
    full.rows_per_world = rows
    e = phys.PhysExecutor(full, lib=emu_lib)
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
        want = o.state()
        for a, b in zip(got, want):
            assert np.array_equal(bits(a[w, :n]), bits(b[0])), f"world {w}"
    # rows beyond the count are untouched
    assert np.array_equal(got[0][3], np.ascontiguousarray(full.pos[3], np.float32))


def test_emu_pair_overflow_reports(emu_lib):
    sc = scenes.tumble(2, num_bodies=40, spread=0.6, seed=5, max_contacts=64)
    e = phys.PhysExecutor(sc, lib=emu_lib)
    with pytest.raises(phys.MadronaError) as ei:
        e.run()
    assert ei.value.code == phys.MDR_EOVERFLOW


def test_emu_rebuild_flag_changes_order_only_when_flagged(port_oracle, emu_lib):
    sc = scenes.tumble(3, num_bodies=20)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    for st in range(10):
        o.step(); e.run()
    o.flag_rebuild(); e.flag_rebuild()
    for st in range(10):
        o.step(); e.run()
        assert_state_equal(e.state(), o.state(), None, f"after rebuild step {st}")


def test_hull_tables_match_oracle(port_oracle, emu_lib):
    sc = scenes.tumble(1)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    for obj in (0, 1):
        a, b = o.get_hull(obj), e.get_hull(obj)
        for k in ("half_edges", "edges", "polys"):
            assert np.array_equal(a[k], b[k]), k
        assert np.array_equal(bits(a["planes"]), bits(b["planes"]))
