"""Generates tests/golden/*.npz from the UNMODIFIED reference compiled by
oracle/Makefile (needs /root/reference). Run in the build container:
`python tests/golden/make_golden.py`. The fixtures pin the oracle restatement
(tests/test_oracle_pin.py) and the CUDA path (tests/test_gpu_parity.py) on
machines where the reference is not mounted.

The values come from oracle/_ref/libmadrona_ref_defined.so: the reference's
sources with automatic variables zero-initialised, so that the manifold slots
buildFaceContactManifold never writes (narrowphase.cpp:830-853) are defined and
EVERY world of every case is comparable. The stock build
(oracle/_ref/libmadrona_ref.so) runs beside it here and must agree bit for bit
in every world that never reads such a slot.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gpu_ecs_madrona_b200 import scenes  # noqa: E402
from oracle import orc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (scene factory, kwargs, snapshot steps, probed world)
    "collisions64": (scenes.collisions64, dict(num_worlds=4), (1, 10, 40, 100), 1),
    "tumble24": (scenes.tumble, dict(num_worlds=6, num_bodies=24), (1, 5, 30), 2),
    "tumble12_noplane": (scenes.tumble, dict(num_worlds=4, num_bodies=12, seed=7, spread=0.8,
                                             with_plane=False), (1, 8), 0),
    # joint constraints (Fixed) beside contacts; hinges only while the reference keeps them finite
    "jointed": (scenes.jointed, dict(num_worlds=6), (1, 10, 60), 3),
    "jointed_hinges": (scenes.jointed, dict(num_worlds=4, hinges=True), (1, 4, 7), 1),
    # non-cube hulls loaded from OBJ files through the asset pipeline (gpu_ecs_madrona_b200/hulls, make_asset_golden.py)
    "hull_pile": (scenes.hull_pile, dict(num_worlds=8), (1, 10, 45), 2),
}


def main():
    orc.build("ref")
    only = set(sys.argv[1:])          # `make_golden.py hull_pile`: (re)generate the named cases only
    for name, (fn, kw, snaps, pw) in CASES.items():
        if only and name not in only:
            continue
        sc = fn(**kw)
        ref = orc.Oracle(sc, "reference_defined")
        # the probe step doubles as step 1
        out = {}
        pr = ref.probe_step(pw)
        out["probe_world"] = np.int32(pw)
        out["probe_pairs"] = pr["pairs"]
        out["probe_num_contacts"] = pr["num_contacts"]
        for s, c in enumerate(pr["contacts"]):
            out[f"probe_contacts_{s}"] = c
        out["probe_slot_boxes"] = pr["slot_boxes"]
        out["probe_leaf_aabbs"] = pr["leaf_aabbs"]
        out["probe_ref_rank"] = pr["ref_rank"]
        # every other world takes its first step too
        rest = orc.Oracle(sc, "reference_defined")
        stock = orc.Oracle(sc, "reference")
        flags = orc.Oracle(sc, "port")      # only for its count of undefined-slot reads per world
        rest.step(); stock.step(); flags.step()
        step = 1
        for snap in snaps:
            while step < snap:
                rest.step(); stock.step(); flags.step()
                step += 1
            pos, rot, vel = rest.state()
            out[f"pos_{snap}"], out[f"rot_{snap}"], out[f"vel_{snap}"] = pos, rot, vel
            det = flags.indeterminate() == 0
            for a, b in zip(rest.state(), stock.state()):
                assert np.array_equal(a.view(np.uint32)[det], b.view(np.uint32)[det]), (name, snap)
        out["undefined_reads"] = flags.indeterminate()
        # the probed world of `ref` must agree with the plain run after step 1
        p1 = ref.state()
        assert np.array_equal(p1[0][pw].view(np.uint32), out["pos_%d" % snaps[0]][pw].view(np.uint32)) \
            or snaps[0] != 1
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, {k: getattr(v, "shape", v) for k, v in out.items() if k.startswith(("pos", "probe_pairs"))})


if __name__ == "__main__":
    main()
