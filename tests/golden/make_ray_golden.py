"""Generates tests/golden/rays_*.npz: BVH::traceRay of the UNMODIFIED reference
(oracle/_ref/libmadrona_ref_defined.so -- the build whose manifold slots are
defined, see make_golden.py; needs /root/reference) on seeded rays, after the
step counts of make_golden.CASES. Run in the build container:
`python tests/golden/make_ray_golden.py`. Pins oracle/port.cpp's restatement and
the CUDA ray kernel where the reference is not mounted.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402
from tests.golden.make_golden import CASES  # noqa: E402
from tests.ray_util import world_rays  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
RAYS = 96
SEED = 23


def main():
    orc.build("ref")
    only = set(sys.argv[1:])          # `make_ray_golden.py hull_pile`: the named cases only
    for name, (fn, kw, snaps, _) in CASES.items():
        if only and name not in only:
            continue
        sc = fn(**kw)
        ref = orc.Oracle(sc, "reference_defined")
        rng = np.random.default_rng(SEED)
        out = {}
        step = 0
        for snap in snaps[:3]:
            ref.step(snap - step)
            step = snap
            o, d = world_rays(rng, RAYS, ref.read_f32(orc.COL_POSITION))
            for tag, tm in (("inf", np.float32(np.inf)), ("t3", np.float32(3.0))):
                res = [ref.trace_rays(w, o[w], d[w], tm) for w in range(sc.num_worlds)]
                out[f"t_{snap}_{tag}"] = np.stack([r[0] for r in res])
                out[f"normal_{snap}_{tag}"] = np.stack([r[1] for r in res])
                out[f"row_{snap}_{tag}"] = np.stack([r[2] for r in res])
            out[f"origins_{snap}"], out[f"dirs_{snap}"] = o, d
        np.savez_compressed(os.path.join(HERE, "rays_" + name + ".npz"), **out)
        print(name, {k: v.shape for k, v in out.items() if k.startswith("row")},
              "hit fraction", float(np.mean([(v >= 0).mean() for k, v in out.items() if k.startswith("row")])))


if __name__ == "__main__":
    main()
