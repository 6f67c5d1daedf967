"""Generates tests/golden/ecs_*.npz from the UNMODIFIED reference ECS + task graph compiled by
oracle/Makefile (oracle/_ref/libmadrona_ecs_ref.so, needs /root/reference). Run in the build
container: `python tests/golden/make_ecs_golden.py`. The fixtures pin the scalar restatement
(tests/test_ecs_oracle_pin.py) and the CUDA path (tests/test_ecs_gpu.py) on machines where the
reference is not mounted. Each fixture is a flat dict "s<step>/<key>" of tests/ecs_util.snapshot.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gpu_ecs_madrona_b200 import ecs_apps as A  # noqa: E402
from tests import ecs_util as U  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# name: (app, worlds, params, snapshot steps)
APP_CASES = {
    "ecs_taskgraph": (A.APP_TASKGRAPH, 3, [12], (1, 40)),
    "ecs_fanout": (A.APP_FANOUT, 2, [24], (1, 6)),
    "ecs_fvs": (A.APP_FVS, 3, [6, 24], (1, 60, 160)),
}
SCRIPT_CASES = {"ecs_script_a": (11, 1, 1500), "ecs_script_b": (12, 0, 1500)}


def script_ops(seed, n):
    return A.random_script(np.random.default_rng(seed), n, max_live=150)


def main():
    for name, (app, W, params, snaps) in APP_CASES.items():
        o = U.make_oracle("reference", app, W, params)
        out, step = {}, 0
        for s in snaps:
            o.step(s - step)
            step = s
            for k, v in U.snapshot(o, U.NUM_EXPORTS[app]).items():
                out[f"s{s}/{k}"] = v
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, len(out), "arrays")
    for name, (seed, single, n) in SCRIPT_CASES.items():
        o = U.make_oracle("reference", A.APP_SCRIPT, 1, [single])
        ops = script_ops(seed, n)
        o.script(0, ops)
        out = {f"s0/{k}": v for k, v in U.snapshot(o).items()}
        out["ops"] = ops
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, len(out), "arrays")


if __name__ == "__main__":
    main()
