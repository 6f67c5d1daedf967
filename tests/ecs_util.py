"""Helpers shared by the ECS parity tests: run an app on a checker and snapshot everything."""
import numpy as np

from gpu_ecs_madrona_b200 import ecs_apps
from oracle import eorc


make_oracle = eorc.make_app_oracle


def snapshot(o, num_exports=0):
    """Everything observable: row counts, every column of every archetype (Entity columns of
    temporaries excluded: they are never written), id maps, exports."""
    snap = {"rows": o.row_counts()}
    for w in range(o.num_worlds):
        for a in range(o.num_archetypes):
            for c in range(len(o.col_bytes[a])):
                snap[f"w{w}_a{a}_c{c}"] = o.read_column(w, a, c)
        ia, ib, gen, st = o.idmap(w)
        snap[f"w{w}_id_a"], snap[f"w{w}_id_b"], snap[f"w{w}_id_gen"], snap[f"w{w}_id_state"] = ia, ib, gen, st
    for s in range(num_exports):
        snap[f"export{s}"] = o.export(s)
    return snap


TEMP_ARCHETYPES = {ecs_apps.APP_SCRIPT: {3}, ecs_apps.APP_TASKGRAPH: set(), ecs_apps.APP_FANOUT: {1},
                   ecs_apps.APP_FVS: {2}}
NUM_EXPORTS = {ecs_apps.APP_SCRIPT: 0, ecs_apps.APP_TASKGRAPH: 2, ecs_apps.APP_FANOUT: 1, ecs_apps.APP_FVS: 2}


def assert_same(a, b, app, what=""):
    assert a.keys() == b.keys()
    for k in a:
        parts = k.split("_")
        if len(parts) == 3 and parts[1][0] == "a" and parts[2] == "c0" and int(parts[1][1:]) in TEMP_ARCHETYPES[app]:
            continue
        assert a[k].shape == b[k].shape, f"{what}{k}: shape {a[k].shape} vs {b[k].shape}"
        assert np.array_equal(a[k], b[k]), f"{what}{k} differs"
