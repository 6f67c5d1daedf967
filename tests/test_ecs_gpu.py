"""GPU parity of the ECS state layer + task-graph executor (CUDA, through the C ABI)
against the checkers of oracle/ecs_api.h: bit-exact row counts, entity ids and generations,
row order, every column, the whole id map (free lists included) and exported columns."""
import numpy as np
import pytest

from gpu_ecs_madrona_b200 import ecs_apps as A
from tests import ecs_util as U
from oracle import eorc

pytestmark = pytest.mark.gpu


def _kind():
    return "reference" if eorc.available("reference") else "port"


def gpu_snapshot(ex, app):
    from gpu_ecs_madrona_b200 import ecs  # noqa: F401
    rows = ex.row_counts()
    snap = {"rows": rows}
    cols = ex.read_world_columns()
    for w in range(ex.num_worlds):
        for a in range(ex.num_archetypes):
            for c in range(len(ex.col_bytes[a])):
                snap[f"w{w}_a{a}_c{c}"] = cols[(a, c)][w, :rows[w, a]]
        ia, ib, gen, st = ex.idmap(w)
        snap[f"w{w}_id_a"], snap[f"w{w}_id_b"], snap[f"w{w}_id_gen"], snap[f"w{w}_id_state"] = ia, ib, gen, st
    for s in range(U.NUM_EXPORTS[app]):
        snap[f"export{s}"] = ex.export(s)
    return snap


def sorted_tmp(snap, app, num_worlds):
    """Temporaries appended by parallel lanes have no defined order unless a sort node ran;
    all our graphs sort them, so nothing to do -- kept as the hook for unsorted graphs."""
    return snap


@pytest.mark.parametrize("app,worlds,steps,every", [
    (A.APP_TASKGRAPH, 37, 60, 20),
    (A.APP_FANOUT, 19, 12, 3),
    (A.APP_FVS, 23, 200, 25),
])
def test_app_parity(app, worlds, steps, every):
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    ex = EcsExecutor(app, worlds)
    ck = U.make_oracle(_kind(), app, worlds)
    U.assert_same(U.snapshot(ck, U.NUM_EXPORTS[app]), gpu_snapshot(ex, app), app, "init ")
    for s in range(steps):
        ex.run()
        ck.step()
        if (s + 1) % every == 0 or s == steps - 1:
            U.assert_same(U.snapshot(ck, U.NUM_EXPORTS[app]), gpu_snapshot(ex, app), app, f"step {s} ")
    ex.close()
    ck.close()


def test_fvs_small_world_sharding_invariance():
    """Worlds [8, 16) of a 16-world run equal a run that owns only those worlds (first_world)."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    full = EcsExecutor(A.APP_FVS, 16, [10, 40])
    part = EcsExecutor(A.APP_FVS, 8, [10, 40], first_world=8)
    full.run(60)
    part.run(60)
    for (a, c), v in part.read_world_columns().items():
        assert np.array_equal(v, full.read_column_all(a, c)[8:]), (a, c)
    assert np.array_equal(part.row_counts(), full.row_counts()[8:])


@pytest.mark.parametrize("seed,with_singleton", [(1, 1), (2, 0), (3, 1), (4, 0)])
def test_script_parity(seed, with_singleton):
    """Random create / destroy / clear / temporary sequences (tests/state.cpp-style churn),
    different per world, replayed in slices: batches of creates and destroys must leave the
    tables and the id map exactly as the reference's one-at-a-time calls do."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    W = 12
    rng = np.random.default_rng(seed)
    ex = EcsExecutor(A.APP_SCRIPT, W, [with_singleton, 2048, 1 << 15])
    ck = U.make_oracle(_kind(), A.APP_SCRIPT, W, [with_singleton])
    scripts = [A.random_script(rng, 3000, max_live=300 + 40 * w) for w in range(W)]
    for k in range(0, 3000, 750):
        ex.script([s[k:k + 750] for s in scripts])
        for w in range(W):
            ck.script(w, scripts[w][k:k + 750])
        U.assert_same(U.snapshot(ck), gpu_snapshot(ex, A.APP_SCRIPT), A.APP_SCRIPT, f"ops {k} ")
    ex.close()
    ck.close()


def test_state_indexing_like_reference_test():
    """tests/state.cpp:38-128 (State.Indexing) at 1/500 scale: create N, destroy every 10th,
    create up to 2N, destroy every 5th -- one world, big table."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    N = 2000
    ops = [(eorc.OP_CREATE, 1, i) for i in range(N)]
    ops += [(eorc.OP_DESTROY, i, 0) for i in range(0, N, 10)]
    ops += [(eorc.OP_CREATE, 1, N + i) for i in range(N)]
    ops += [(eorc.OP_DESTROY, i, 0) for i in range(0, 2 * N, 5)]
    ops = np.array(ops, eorc.OP_DTYPE)
    ex = EcsExecutor(A.APP_SCRIPT, 1, [1, 4096, 1 << 13])
    ck = U.make_oracle(_kind(), A.APP_SCRIPT, 1, [1])
    ex.script([ops])
    ck.script(0, ops)
    U.assert_same(U.snapshot(ck), gpu_snapshot(ex, A.APP_SCRIPT), A.APP_SCRIPT)


def test_state_indexing_at_the_reference_scale():
    """tests/state.cpp:38-128 (State.Indexing) at its own size: 1 000 000 entities created, every 10th destroyed,
    refilled to 2 000 000, every 5th destroyed -- one world, every id, generation, row and free-list link
    against the checker."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    N = 1_000_000
    ops = np.zeros(N + N // 10 + N + N // 10 + (2 * N) // 5, eorc.OP_DTYPE)
    k = 0
    def put(op, a, b):
        nonlocal k
        n = len(a)
        ops["op"][k:k + n], ops["a"][k:k + n], ops["b"][k:k + n] = op, a, b
        k += n
    ar = np.arange
    put(eorc.OP_CREATE, np.ones(N, np.int32), ar(N))
    put(eorc.OP_DESTROY, ar(0, N, 10), 0)
    put(eorc.OP_CREATE, np.ones(N + N // 10, np.int32), N + ar(N + N // 10))
    put(eorc.OP_DESTROY, ar(0, 2 * N + N // 10, 5)[: (2 * N) // 5], 0)
    assert k == len(ops)
    cap = 2 * N + 64
    ex = EcsExecutor(A.APP_SCRIPT, 1, [1, cap, 2 * N + N // 10 + 64])
    ck = U.make_oracle(_kind(), A.APP_SCRIPT, 1, [1])
    ex.script([ops])
    ck.script(0, ops)
    # 2.1 N handles made, N / 10 destroyed, then 2 N / 5 destroys of which N / 10 name handles already dead
    assert ex.row_counts()[0, 1] == 2 * N - (2 * N) // 5 + N // 10 == ck.row_counts()[0, 1]
    U.assert_same(U.snapshot(ck), gpu_snapshot(ex, A.APP_SCRIPT), A.APP_SCRIPT)
    ex.close()
    ck.close()


def test_capacity_overflow_is_reported():
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    from gpu_ecs_madrona_b200 import MadronaError
    ex = EcsExecutor(A.APP_SCRIPT, 2, [0, 16, 64])
    ops = np.array([(eorc.OP_CREATE, 0, i) for i in range(40)], eorc.OP_DTYPE)
    with pytest.raises(MadronaError) as ei:
        ex.script([ops, ops[:4]])
    assert ei.value.code == -4
    assert list(ex.row_counts()[:, 0]) == [16, 4]


def test_registry_rejects_bad_arguments():
    from gpu_ecs_madrona_b200.ecs import EcsRegistry
    from gpu_ecs_madrona_b200 import MadronaError
    r = EcsRegistry(4)
    with pytest.raises(MadronaError):
        r.register_component(6)            # not a multiple of 4
    c = r.register_component(12)
    with pytest.raises(MadronaError):
        r.register_archetype([c + 5], 10)  # unknown component
    a = r.register_archetype([c], 10)
    with pytest.raises(MadronaError):
        r.parallel_for(1, [c], deps=[3])   # dependency on a node that does not exist yet
    with pytest.raises(MadronaError):
        r.run()                            # not built
    r.build()
    with pytest.raises(MadronaError):
        r.register_component(4)            # after build
    assert a == 0


@pytest.mark.parametrize("name", ["ecs_taskgraph", "ecs_fanout", "ecs_fvs"])
def test_gpu_matches_golden(name):
    """The committed fixtures were generated from the compiled reference (tests/golden/make_ecs_golden.py)."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    from tests.golden import make_ecs_golden as G
    from tests.util import load_golden
    app, W, params, snaps = G.APP_CASES[name]
    g = load_golden(name)
    ex = EcsExecutor(app, W, params)
    step = 0
    for s in snaps:
        ex.run(s - step)
        step = s
        pre = f"s{s}/"
        want = {k[len(pre):]: g[k] for k in g.files if k.startswith(pre)}
        U.assert_same(want, gpu_snapshot(ex, app), app, f"{name} step {s} ")


def test_full_size_fvs_properties():
    """BASELINE.json configs[3] size (8192 worlds x 250 units): size-independent properties --
    row totals conserved by respawning, every live entity's Loc points back at its row, ids unique."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    ex = EcsExecutor(A.APP_FVS, 8192)
    ex.run(150)
    rows = ex.row_counts()
    assert (rows[:, 0] == 50).all() and (rows[:, 1] == 200).all() and (rows[:, 2] == 0).all()
    ents = {a: ex.read_column_all(a, 0).view(np.int32).reshape(8192, -1, 2) for a in (0, 1)}
    churn = 0
    for w in (0, 1, 4095, 8191):
        ia, ib, gen, st = ex.idmap(w)
        seen = set()
        for a in (0, 1):
            for r in range(rows[w, a]):
                g, i = ents[a][w, r]
                assert gen[i] == g and ia[i] == a and ib[i] == r
                assert i not in seen
                seen.add(i)
        churn += int(gen.sum())
    assert churn > 0
    tick = ex.read_column_all(3, 1).view(np.uint32).reshape(8192, -1)[:, 0]
    assert (tick == 150).all()


def test_zero_copy_tensor_views():
    """getExported -> torch (the reference's py::Tensor role): views alias the tables, no copy."""
    import torch
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    from gpu_ecs_madrona_b200 import phys, scenes
    ex = EcsExecutor(A.APP_TASKGRAPH, 5)
    t = ex.export_tensor(0)                       # Agent Position, staged export: [5 rows, 3]
    assert t.shape == (5, 3) and t.is_cuda
    assert np.array_equal(t.cpu().numpy().view(np.uint8).ravel(), ex.export(0))
    col = ex.column_tensor(0, 1)                  # Sphere Position column, in place
    before = col[:, :100].clone()
    ex.run(3)
    torch.cuda.synchronize()
    assert not torch.equal(before, col[:, :100])  # the view sees the step's writes
    assert np.array_equal(col.cpu().numpy().view(np.uint8).reshape(5, -1, 12), ex.read_column_all(0, 1))
    pe = phys.PhysExecutor(scenes.collisions64(3))
    pos = pe.column_tensor(phys.COL_POSITION)
    pe.run(2)
    assert np.array_equal(pos.cpu().numpy()[:, :65], pe.state()[0])


@pytest.mark.parametrize("app,worlds,steps,sample", [
    (A.APP_TASKGRAPH, 1024, 1000, [0, 1, 77, 500, 1023]),          # BASELINE.json configs[0] at its own size
    (A.APP_FANOUT, 4096, 25, [0, 9, 2048, 4095]),                  # configs[1]
    (A.APP_FVS, 8192, 150, [0, 3, 1234, 4096, 8190, 8191]),        # configs[3]
])
def test_full_size_sampled_worlds_match_the_checker(app, worlds, steps, sample):
    """The whole batch of a BASELINE config on the GPU; sampled worlds against the checker run on those worlds
    alone (worlds are independent and seeded by their GLOBAL index, so a one-world checker with first_world = w
    is world w of the batch): row counts, Entity columns, every component, the id map -- bit for bit."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    ex = EcsExecutor(app, worlds)
    ex.run(steps)
    rows = ex.row_counts()
    cols = ex.read_world_columns()
    temp = U.TEMP_ARCHETYPES[app]
    for w in sample:
        ck = U.make_oracle(_kind(), app, 1, None, first_world=w)
        ck.step(steps)
        assert np.array_equal(ck.row_counts()[0], rows[w]), w
        for a in range(ex.num_archetypes):
            for c in range(len(ex.col_bytes[a])):
                if c == 0 and a in temp:
                    continue
                assert np.array_equal(cols[(a, c)][w, :rows[w, a]], ck.read_column(0, a, c)), (w, a, c)
        for x, y in zip(ex.idmap(w), ck.idmap(0)):
            assert np.array_equal(x, y), w
        ck.close()
    ex.close()


def test_forty_node_graph():
    """The reference's own physics graph is 40 nodes at 4 substeps: a 40-node graph here --
    20 x (drift, clamp) over the simple_taskgraph tables -- equals 20 steps of the 2-node world on the checker."""
    from gpu_ecs_madrona_b200.ecs import EcsRegistry
    from gpu_ecs_madrona_b200 import ecs_apps
    W, n = 9, 100
    r = EcsRegistry(W)
    pos, rot = r.register_component(12), r.register_component(16)
    sphere = r.register_archetype([pos, rot], n)
    agent = r.register_archetype([pos, rot], 1)
    r.export_column(agent, pos, 0)
    r.export_column(agent, rot, 1)
    prev = []
    for k in range(20):
        d = r.parallel_for(1, [pos], deps=prev)          # kTgDrift
        prev = [r.parallel_for(2, [pos], deps=[d])]      # kTgClamp
    r.build()
    r.make_entities(sphere, n)
    r.make_entities(agent, 1)
    for (a, c), data in ecs_apps.initial_columns(A.APP_TASKGRAPH, W, [n]).items():
        r.write_column_all(a, c, data)
    r.run(1)
    ck = U.make_oracle(_kind(), A.APP_TASKGRAPH, W, [n])
    ck.step(20)
    for a, rows in ((0, n), (1, 1)):
        got = r.read_column_all(a, 1, rows)
        for w in range(W):
            assert np.array_equal(got[w], ck.read_column(w, a, 1)), (a, w)
    assert np.array_equal(r.export(0), ck.export(0))
    r.close()
    ck.close()


def test_one_launch_step_equals_the_four_launch_graph(monkeypatch):
    """The fused step (copy-in, nodes, grid barrier, offsets, copy-out in one kernel; used when every world is
    resident at once) against the run graph of four launches (MDR_ECS_NO_FUSED=1; also what larger batches use):
    every column, row counts, id maps and the staged exports with their offsets."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    monkeypatch.setenv("MDR_ECS_FUSED", "1")            # on by default from 2048 worlds
    fused = EcsExecutor(A.APP_FVS, 300, [10, 40])
    monkeypatch.delenv("MDR_ECS_FUSED")
    monkeypatch.setenv("MDR_ECS_NO_FUSED", "1")
    plain = EcsExecutor(A.APP_FVS, 300, [10, 40])
    monkeypatch.delenv("MDR_ECS_NO_FUSED")
    assert fused.launches_per_step() >= 1
    for steps in (1, 7, 60):
        fused.run(steps)
        plain.run(steps)
        assert fused.launches_per_step() == 1 and plain.launches_per_step() == 4
        assert np.array_equal(fused.row_counts(), plain.row_counts())
        for (a, c), v in fused.read_world_columns().items():
            if c == 0 and a in U.TEMP_ARCHETYPES[A.APP_FVS]:
                continue
            rows = fused.row_counts()[:, a]
            w = plain.read_column_all(a, c)
            for k in (0, 1, 150, 299):
                assert np.array_equal(v[k, :rows[k]], w[k, :rows[k]]), (steps, a, c, k)
        for s in range(2):
            assert np.array_equal(fused.export(s), plain.export(s)), (steps, s)
    fused.close()
    plain.close()


@pytest.mark.parametrize("app,worlds,params", [
    (A.APP_TASKGRAPH, 70, None),
    (A.APP_FANOUT, 21, None),
    (A.APP_FVS, 300, [10, 40]),
])
@pytest.mark.parametrize("no_fused", [False, True])
def test_batched_steps_equal_single_steps(monkeypatch, app, worlds, params, no_fused):
    """run(n > 1) walks a world's graph n times inside one launch (staged exports copied in once before, offsets and
    copy-out once after). Against n single-step graph launches (MDR_ECS_NO_BATCH=1): every column, row counts, id
    maps, staged exports -- with the one-launch kernel and with the four-launch form (MDR_ECS_NO_FUSED=1)."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    args = (app, worlds) if params is None else (app, worlds, params)
    if no_fused:
        monkeypatch.setenv("MDR_ECS_NO_FUSED", "1")
    batch = EcsExecutor(*args)
    monkeypatch.setenv("MDR_ECS_NO_BATCH", "1")
    single = EcsExecutor(*args)
    monkeypatch.delenv("MDR_ECS_NO_BATCH")
    staged = single.launches_per_step() > 1 or batch.launches_per_step() > 1
    for steps in (5, 1, 33):
        batch.run(steps)
        single.run(steps)
        if steps > 1:
            assert batch.last_launches() in (1, 4) and batch.last_launches() == (4 if staged and no_fused else 1)
            assert single.last_launches() == steps * single.launches_per_step()
        U.assert_same(gpu_snapshot(batch, app), gpu_snapshot(single, app), app, f"after run({steps}) ")
    batch.close()
    single.close()


@pytest.mark.parametrize("app,worlds", [(A.APP_TASKGRAPH, 50), (A.APP_FVS, 40)])
def test_step_read_equals_step_plus_exports(app, worlds):
    """mdr_ecs_step_read (one step and the exported columns in host memory, two synchronisations) against
    run(1) + export(slot): the same bytes, step after step."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    a, b = EcsExecutor(app, worlds), EcsExecutor(app, worlds)
    n = U.NUM_EXPORTS[app]
    bufs = [a.pinned_export_buffer(s) for s in range(n)]
    for step in range(12):
        got = a.step_read(list(range(n)), bufs)
        b.run(1)
        for s in range(n):
            want = b.export(s)
            assert got[s] == len(want), (step, s)
            assert np.array_equal(bufs[s].reshape(-1).view(np.uint8)[:got[s]], want), (step, s)
    a.close()
    b.close()
