"""BVH::traceRay (src/physics/broadphase.cpp:653-856) parity: the restatement in
oracle/port.cpp against the compiled reference and its golden fixtures
(tests/golden/make_ray_golden.py), the device source compiled for the host
(tests/emu) against the restatement, and -- on the GPU -- the CUDA kernel
through mdr_phys_trace_rays against both. Bit-exact: hit distance, normal,
hit row."""
import numpy as np
import pytest

from gpu_ecs_madrona_b200 import phys, scenes
from tests.golden.make_golden import CASES
from tests.golden.make_ray_golden import RAYS, SEED
from tests.ray_util import assert_rays_equal, world_rays
from tests.util import load_golden

SCENES = {
    "collisions64": lambda: scenes.collisions64(6),
    "collisions16": lambda: scenes.collisions64(5, num_cubes=16),
    "tumble24": lambda: scenes.tumble(12, seed=5),
    "tumble12_noplane": lambda: scenes.tumble(6, num_bodies=12, seed=7, spread=0.8, with_plane=False),
}


def oracle_rays(o, origins, dirs, t_max):
    res = [o.trace_rays(w, origins[w], dirs[w], t_max) for w in range(origins.shape[0])]
    return tuple(np.stack([r[k] for r in res]) for k in range(3))


def golden_walk(name, executor_rays, step_fn):
    """Replays make_ray_golden's schedule and compares at every snapshot, every world
    (the fixtures come from the reference build whose manifold slots are defined)."""
    fn, kw, snaps, _ = CASES[name]
    g = load_golden("rays_" + name)
    step = 0
    for snap in snaps[:3]:
        step_fn(snap - step)
        step = snap
        o, d = g[f"origins_{snap}"], g[f"dirs_{snap}"]
        worlds = np.arange(o.shape[0])
        for tag, tm in (("inf", np.float32(np.inf)), ("t3", np.float32(3.0))):
            want = (g[f"t_{snap}_{tag}"], g[f"normal_{snap}_{tag}"], g[f"row_{snap}_{tag}"])
            assert_rays_equal(executor_rays(o, d, tm), want, worlds, f"{name} step {snap} {tag}")


@pytest.mark.parametrize("name", sorted(CASES))
def test_port_rays_match_reference_golden(port_oracle, name):
    fn, kw, _, _ = CASES[name]
    o = port_oracle.Oracle(fn(**kw), "port")
    golden_walk(name, lambda a, b, tm: oracle_rays(o, a, b, tm), o.step)


def test_golden_ray_generator_is_seeded_like_the_tests():
    g = load_golden("rays_collisions64")
    assert g["origins_1"].shape[1] == RAYS and SEED == 23


@pytest.mark.parametrize("name", sorted(SCENES))
def test_port_rays_match_compiled_reference(port_oracle, ref_oracle, name):
    sc = SCENES[name]()
    ref, port = ref_oracle.Oracle(sc, "reference_defined"), port_oracle.Oracle(sc, "port")
    rng = np.random.default_rng(11)
    for steps in (1, 6, 20):
        ref.step(steps)
        port.step(steps)
        ok = np.arange(sc.num_worlds)
        o, d = world_rays(rng, 128, ref.read_f32(ref_oracle.COL_POSITION))
        for tm in (np.float32(np.inf), np.float32(2.5)):
            assert_rays_equal(oracle_rays(port, o, d, tm), oracle_rays(ref, o, d, tm), ok, f"{name} +{steps}")


@pytest.mark.parametrize("name", sorted(SCENES))
def test_emu_rays_match_oracle(port_oracle, emu_lib, name):
    sc = SCENES[name]()
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    rng = np.random.default_rng(3)
    for steps in (1, 7, 22):
        o.step(steps)
        e.run(steps)
        ok = np.arange(sc.num_worlds)
        og, dr = world_rays(rng, 64, o.read_f32(port_oracle.COL_POSITION))
        for tm in (None, np.float32(2.0)):
            t, n, ent, row = e.trace_rays(og, dr, tm)
            want = oracle_rays(o, og, dr, np.float32(np.inf) if tm is None else tm)
            assert_rays_equal((t, n, row), want, ok, f"{name} +{steps}")
            ids = e.read(phys.COL_ENTITY).view(np.int32)
            hit = row >= 0
            assert np.array_equal(ent[hit], ids[np.nonzero(hit)[0], row[hit]])
            assert (ent[~hit] == -1).all()


# ------------------------------------------------------------------- GPU ----
@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(SCENES))
def test_cuda_rays_match_oracle(cuda_lib, port_oracle, name):
    sc = SCENES[name]()
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc)
    rng = np.random.default_rng(3)
    for steps, rays in ((1, 33), (7, 128), (22, 300)):
        o.step(steps)
        e.run(steps)
        ok = np.arange(sc.num_worlds)
        og, dr = world_rays(rng, rays, o.read_f32(port_oracle.COL_POSITION))
        for tm in (None, np.float32(2.0)):
            t, n, ent, row = e.trace_rays(og, dr, tm)
            want = oracle_rays(o, og, dr, np.float32(np.inf) if tm is None else tm)
            assert_rays_equal((t, n, row), want, ok, f"{name} +{steps}")
            ids = e.read(phys.COL_ENTITY).view(np.int32)
            hit = row >= 0
            assert np.array_equal(ent[hit], ids[np.nonzero(hit)[0], row[hit]])
            assert (ent[~hit] == -1).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_rays_match_reference_golden(cuda_lib, port_oracle, name):
    fn, kw, _, _ = CASES[name]
    e = phys.PhysExecutor(fn(**kw))

    def rays(og, d, tm):
        t, n, _, row = e.trace_rays(og, d, tm)
        return t, n, row

    golden_walk(name, rays, e.run)


@pytest.mark.gpu
def test_cuda_rays_ragged_worlds_and_device_pointers(cuda_lib, port_oracle):
    """Worlds with fewer rows than the capacity (one empty); the _device entry on torch tensors."""
    import torch
    full = scenes.tumble(5, num_bodies=15)
    rows = np.array([16, 9, 1, 0, 5], np.int32)
    full.rows_per_world = rows
    e = phys.PhysExecutor(full)
    e.run(3)
    rng = np.random.default_rng(9)
    pos = e.state()[0]
    og, dr = world_rays(rng, 40, np.where(np.isfinite(pos), pos, 0))
    t, n, ent, row = e.trace_rays(og, dr)
    for w in range(5):
        assert (row[w] < max(rows[w], 1)).all()
    assert (row[3] == -1).all() and np.isinf(t[3]).all()
    dev = [torch.from_numpy(a).cuda() for a in (og, dr)]
    out_t = torch.zeros(5, 40, device="cuda")
    out_n = torch.zeros(5, 40, 3, device="cuda")
    out_row = torch.zeros(5, 40, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = e.lib.mdr_phys_trace_rays_device(e.h, 40, dev[0].data_ptr(), dev[1].data_ptr(), None,
                                          out_t.data_ptr(), out_n.data_ptr(), None, out_row.data_ptr())
    assert rc == 0
    e.sync()
    assert np.array_equal(out_row.cpu().numpy(), row)
    assert np.array_equal(out_t.cpu().numpy().view(np.uint32), t.view(np.uint32))
