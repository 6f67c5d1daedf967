"""Pins the ECS checkers (CPU only): the scalar restatement oracle/ecs_port.cpp must equal
(a) the golden fixtures generated from the compiled reference and (b), where the compiled
reference is present, the reference itself on fresh runs -- bit for bit on row counts, Entity
columns, every component column, the id map with its free lists, and exported columns.
Also re-expresses the reference's own ECS tests that apply to this path
(tests/state.cpp:38-266, tests/id_map.cpp:56-89) as op scripts."""
import numpy as np
import pytest

from gpu_ecs_madrona_b200 import ecs_apps as A
from oracle import eorc
from tests import ecs_util as U
from tests.golden import make_ecs_golden as G
from tests.util import load_golden


@pytest.fixture(scope="module", autouse=True)
def built():
    import subprocess, os
    subprocess.run(["make", "-C", os.path.dirname(eorc.HERE + "/"), "libmadrona_ecs_port.so"], check=True,
                   capture_output=True)


def golden_step(g, s):
    pre = f"s{s}/"
    return {k[len(pre):]: g[k] for k in g.files if k.startswith(pre)}


@pytest.mark.parametrize("name", sorted(G.APP_CASES))
def test_port_matches_golden_apps(name):
    app, W, params, snaps = G.APP_CASES[name]
    g = load_golden(name)
    o = U.make_oracle("port", app, W, params)
    step = 0
    for s in snaps:
        o.step(s - step)
        step = s
        U.assert_same(golden_step(g, s), U.snapshot(o, U.NUM_EXPORTS[app]), app, f"{name} step {s} ")


@pytest.mark.parametrize("name", sorted(G.SCRIPT_CASES))
def test_port_matches_golden_scripts(name):
    seed, single, n = G.SCRIPT_CASES[name]
    g = load_golden(name)
    o = U.make_oracle("port", A.APP_SCRIPT, 1, [single])
    o.script(0, g["ops"])
    U.assert_same(golden_step(g, 0), U.snapshot(o), A.APP_SCRIPT, name)


needs_ref = pytest.mark.skipif(not eorc.available("reference"), reason="compiled reference not on this machine")


@needs_ref
@pytest.mark.parametrize("app,steps", [(A.APP_TASKGRAPH, 30), (A.APP_FANOUT, 6), (A.APP_FVS, 120)])
def test_port_matches_reference_apps(app, steps):
    r, p = U.make_oracle("reference", app, 4), U.make_oracle("port", app, 4)
    for s in range(steps):
        r.step()
        p.step()
        if s % 15 == 0 or s == steps - 1:
            U.assert_same(U.snapshot(r, U.NUM_EXPORTS[app]), U.snapshot(p, U.NUM_EXPORTS[app]), app, f"step {s} ")


@needs_ref
@pytest.mark.parametrize("seed", [21, 22, 23, 24])
def test_port_matches_reference_scripts(seed):
    rng = np.random.default_rng(seed)
    r, p = U.make_oracle("reference", A.APP_SCRIPT, 1, [seed % 2]), U.make_oracle("port", A.APP_SCRIPT, 1, [seed % 2])
    ops = A.random_script(rng, 5000, max_live=500, p_clear=0.004)
    for k in range(0, len(ops), 1000):
        r.script(0, ops[k:k + 1000])
        p.script(0, ops[k:k + 1000])
        U.assert_same(U.snapshot(r), U.snapshot(p), A.APP_SCRIPT, f"ops {k} ")


def _kinds():
    return ["port"] + (["reference"] if eorc.available("reference") else [])


@pytest.mark.parametrize("kind", _kinds())
def test_state_indexing(kind):
    """tests/state.cpp:38-128 (State.Indexing) at reduced size: handles stay valid and values
    follow their entity through swap-removes."""
    N = 5000
    o = U.make_oracle(kind, A.APP_SCRIPT, 1, [0])
    o.script(0, np.array([(eorc.OP_CREATE, 0, i) for i in range(N)], eorc.OP_DTYPE))
    o.script(0, np.array([(eorc.OP_DESTROY, i, 0) for i in range(0, N, 10)], eorc.OP_DTYPE))
    assert o.num_rows(0, 0) == N - N // 10
    ents = o.read_column(0, 0, 0).view(np.int32).reshape(-1, 2)
    vals = o.read_column(0, 0, 1).view(np.uint32).ravel()
    a, b, gen, _ = o.idmap(0)
    for row in range(len(ents)):
        g, i = ents[row]
        assert gen[i] == g and a[i] == 0 and b[i] == row     # Loc of a live entity points at its row
        assert vals[row] == i and vals[row] % 10 != 0        # ids were handed out densely from 0
    o.script(0, np.array([(eorc.OP_CREATE, 0, N + i) for i in range(N)], eorc.OP_DTYPE))
    assert o.num_rows(0, 0) == 2 * N - N // 10
    o.script(0, np.array([(eorc.OP_DESTROY, i, 0) for i in range(0, 2 * N, 5)], eorc.OP_DTYPE))
    ents = o.read_column(0, 0, 0).view(np.int32).reshape(-1, 2)
    a, b, gen, _ = o.idmap(0)
    assert all(gen[i] == g and b[i] == r for r, (g, i) in enumerate(ents))


@pytest.mark.parametrize("kind", _kinds())
def test_state_reset_invalidates_handles(kind):
    """tests/state.cpp:227-266 (State.Reset): clear<Archetype> frees every id; old handles
    no longer resolve, new entities do."""
    o = U.make_oracle(kind, A.APP_SCRIPT, 1, [0])
    o.script(0, np.array([(eorc.OP_CREATE, 2, i) for i in range(200)], eorc.OP_DTYPE))
    before = o.read_column(0, 2, 0).view(np.int32).reshape(-1, 2).copy()
    o.script(0, np.array([(eorc.OP_CLEAR, 2, 0)], eorc.OP_DTYPE))
    assert o.num_rows(0, 2) == 0
    _, _, gen, _ = o.idmap(0)
    assert all(gen[i] != g for g, i in before)
    o.script(0, np.array([(eorc.OP_DESTROY, i, 0) for i in range(200)], eorc.OP_DTYPE))   # stale: no-ops
    o.script(0, np.array([(eorc.OP_CREATE, 2, 7)] * 10, eorc.OP_DTYPE))
    assert o.num_rows(0, 2) == 10


@pytest.mark.parametrize("kind", _kinds())
def test_id_map_acquire_release_cycles(kind):
    """tests/id_map.cpp:56-89 (IDs.Threads) restated for one world: acquire 1000, all present,
    release, repeated; the id space must not grow after the first round."""
    o = U.make_oracle(kind, A.APP_SCRIPT, 1, [0])
    made = 0
    sizes = []
    for it in range(6):
        o.script(0, np.array([(eorc.OP_CREATE, 0, i) for i in range(1000)], eorc.OP_DTYPE))
        ents = o.read_column(0, 0, 0).view(np.int32).reshape(-1, 2)
        _, _, gen, _ = o.idmap(0)
        assert len(ents) == 1000 and len(set(ents[:, 1])) == 1000
        assert all(gen[i] == g for g, i in ents)
        o.script(0, np.array([(eorc.OP_DESTROY, made + i, 0) for i in range(1000)], eorc.OP_DTYPE))
        made += 1000
        sizes.append(len(gen))
    assert sizes[-1] == sizes[0] == 1024


def test_multi_column_alignment_and_values():
    """tests/state.cpp:130-190 (State.MultiColumn): 1001 entities x 3 components incl. a
    16-byte aligned one; swap-remove keeps every entity's values together."""
    o = U.make_oracle("port", A.APP_SCRIPT, 1, [0])
    o.script(0, np.array([(eorc.OP_CREATE, 2, i) for i in range(1001)], eorc.OP_DTYPE))
    o.script(0, np.array([(eorc.OP_DESTROY, i, 0) for i in range(0, 1001, 3)], eorc.OP_DTYPE))
    va = o.read_column(0, 2, 1).view(np.uint32).ravel()
    vb = o.read_column(0, 2, 2).view(np.float32).reshape(-1, 3)
    vc = o.read_column(0, 2, 3).view(np.float32).reshape(-1, 4)
    assert o.col_bytes[2] == [8, 4, 12, 16]
    assert np.array_equal(vb[:, 0], va.astype(np.float32)) and np.array_equal(vc[:, 0], va.astype(np.float32) * 2)
    assert not (va % 3 == 0).any()
