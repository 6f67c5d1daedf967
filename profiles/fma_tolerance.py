#!/usr/bin/env python
"""what the FMA-contracted build changes. For each scene: the exact build, the FMA build
and the oracle from the same initial state; after 1 step and after 100 steps the largest
|fma - oracle| / max(|oracle|, 1) over positions / rotations / velocities, whether candidate-pair
counts and contact counts per world are identical, and the probed pair list of world 0.

  python profiles/fma_tolerance.py [--exact LIB] [--fma LIB]
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel_err(a, b):
    a = a.astype(np.float64)
    b = b.astype(np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--exact", default="libmadrona_b200.so")
    ap.add_argument("--fma", default="libmadrona_b200_fma.so")
    a = ap.parse_args()
    from gpu_ecs_madrona_b200 import phys, scenes
    from oracle import orc
    libs = {k: phys.bind(C.CDLL(os.path.join(ROOT, "gpu_ecs_madrona_b200", v))) for k, v in (("exact", a.exact), ("fma", a.fma))}
    cases = {
        "collisions64 (64 worlds)": lambda: scenes.collisions64(64),
        "tumble (48 worlds x 24 boxes)": lambda: scenes.tumble(48),
        "hull_pile (32 worlds x 20 hulls)": lambda: scenes.hull_pile(32),
    }
    print("| scene | step | exact == oracle (bits) | fma vs oracle: pos | rot | vel | pair counts equal | contact counts equal | pair list world 0 equal |")
    print("|---|---|---|---|---|---|---|---|---|")
    for name, mk in cases.items():
        sc = mk()
        o = orc.Oracle(sc, "port")
        ex = phys.PhysExecutor(sc, lib=libs["exact"])
        fm = phys.PhysExecutor(sc, lib=libs["fma"])
        done = 0
        for upto in (1, 100):
            n = upto - done
            if upto == 1:
                pe, pf = ex.probe_step(0), fm.probe_step(0)
                pairs_eq = np.array_equal(pe["pairs"], pf["pairs"])
            else:
                ex.run(n); fm.run(n)
                pairs_eq = None
            o.step(n, 8)
            done = upto
            se, sf, so = ex.state(), fm.state(), o.state()
            exact_ok = all(np.array_equal(x.view(np.uint32), y.view(np.uint32)) for x, y in zip(se, so))
            errs = [rel_err(x, y) for x, y in zip(sf, so)]
            (pe_, ce_), (pf_, cf_) = ex.stats(), fm.stats()
            print(f"| {name} | {upto} | {exact_ok} | {errs[0]:.2e} | {errs[1]:.2e} | {errs[2]:.2e} | "
                  f"{np.array_equal(pe_, pf_)} | {np.array_equal(ce_, cf_)} | {pairs_eq if pairs_eq is not None else '-'} |", flush=True)
        ex.close(); fm.close(); o.close()


if __name__ == "__main__":
    main()
