"""The observation feeder (instanceTransformSetup / updateViewData,
src/mw/render/interop.cpp:10-96) against the reference's own inline arithmetic
(oracle/ref_harness.cpp calls Mat3x4::fromTRS of the reference headers), the
restatement, the host build of the device source and -- on the GPU -- the CUDA
kernels through mdr_phys_pack_instances / mdr_phys_pack_views. Bit-exact."""
import numpy as np
import pytest

from gpu_ecs_madrona_b200 import phys, scenes

OFFSET = np.array([3.5, -2.25, 0.125], np.float32)
V = 3


def view_spec(rng, W, N):
    rows = rng.integers(0, N, (W, V)).astype(np.int32)
    cam = rng.uniform(-1, 1, (W, V, 3)).astype(np.float32)
    ys = rng.uniform(0.3, 1.2, (W, V)).astype(np.float32)
    return rows, cam, ys


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def test_port_feeder_matches_compiled_reference(port_oracle, ref_oracle):
    sc = scenes.tumble(5, seed=2)
    # the reference build whose manifold slots are defined (oracle/Makefile): every world
    ref, port = ref_oracle.Oracle(sc, "reference_defined"), port_oracle.Oracle(sc, "port")
    rng = np.random.default_rng(1)
    for steps in (0, 3, 9):
        ref.step(steps)
        port.step(steps)
        spec = view_spec(rng, sc.num_worlds, sc.bodies_per_world)
        a, b = ref.feeder(OFFSET, *spec), port.feeder(OFFSET, *spec)
        for x, y in zip(a, b):
            assert np.array_equal(bits(x), bits(y))


def run_feeder_host(e, spec):
    W, st = e.W, e.row_stride()
    tr = np.full((W, st, 12), np.nan, np.float32)
    ids = np.zeros((W, st), np.int32)
    views = np.zeros((W, V, 8), np.float32)
    e.pack_instances(tr.ctypes.data, ids.ctypes.data, OFFSET)
    e.pack_views(spec[0].ctypes.data, spec[1].ctypes.data, spec[2].ctypes.data, views.ctypes.data, V, OFFSET)
    return tr, ids, views


def check(e, o, sc, tr, ids, views, spec):
    N = sc.bodies_per_world
    want_tr, want_views = o.feeder(OFFSET, *spec)
    assert np.array_equal(bits(tr[:, :N]), bits(want_tr))
    assert np.array_equal(bits(views), bits(want_views))
    assert np.array_equal(ids[:, :N], np.asarray(sc.obj_ids, np.int32).reshape(sc.num_worlds, N))
    assert (ids[:, N:] == -1).all() and not tr[:, N:].any()


def test_emu_feeder_matches_oracle(port_oracle, emu_lib):
    sc = scenes.tumble(6, seed=4)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc, lib=emu_lib)
    rng = np.random.default_rng(7)
    for steps in (1, 5):
        o.step(steps)
        e.run(steps)
        spec = view_spec(rng, sc.num_worlds, sc.bodies_per_world)
        check(e, o, sc, *run_feeder_host(e, spec), spec)


@pytest.mark.gpu
def test_cuda_feeder_matches_oracle(cuda_lib, port_oracle):
    import torch
    sc = scenes.tumble(40, seed=4)
    o = port_oracle.Oracle(sc, "port")
    e = phys.PhysExecutor(sc)
    rng = np.random.default_rng(7)
    W, st = sc.num_worlds, e.row_stride()
    for steps in (1, 5):
        o.step(steps)
        e.run(steps)
        spec = view_spec(rng, W, sc.bodies_per_world)
        dev = [torch.from_numpy(a).cuda() for a in spec]
        tr = torch.full((W, st, 12), float("nan"), device="cuda")
        ids = torch.zeros(W, st, dtype=torch.int32, device="cuda")
        views = torch.zeros(W, V, 8, device="cuda")
        torch.cuda.synchronize()
        e.pack_instances(tr.data_ptr(), ids.data_ptr(), OFFSET)
        e.pack_views(dev[0].data_ptr(), dev[1].data_ptr(), dev[2].data_ptr(), views.data_ptr(), V, OFFSET)
        e.sync()
        check(e, o, sc, tr.cpu().numpy(), ids.cpu().numpy(), views.cpu().numpy(), spec)
