#!/usr/bin/env python
"""Where the end-to-end step (mdr_phys_step_io_async, collisions-64, 16384 worlds) spends its time: the same call with
no copies, inputs only, outputs only, both; synchronised after every step (the bench's protocol) or once at the end.
python profiles/io_probe.py [--chunks 14]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_ecs_madrona_b200 import phys, scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--chunks", type=int, default=14)
ap.add_argument("--steps", type=int, default=30)
a = ap.parse_args()
sc = scenes.collisions64(16384, max_contacts=640)
ex = phys.PhysExecutor(sc)
ex.run(130)
ins = [(c, ex.pinned(c)) for c in (phys.COL_EXT_FORCE, phys.COL_EXT_TORQUE)]
outs = [(c, ex.pinned(c)) for c in (phys.COL_POSITION, phys.COL_ROTATION, phys.COL_VELOCITY)]
for _, arr in ins:
    arr[:] = 0
for name, i, o in (("no copies", [], []), ("inputs only", ins, []), ("outputs only", [], outs), ("both", ins, outs)):
    for per_step_sync in (True, False):
        for _ in range(3):
            ex.step_io_async(i, o, a.chunks)
        ex.sync()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            ex.step_io_async(i, o, a.chunks)
            if per_step_sync:
                ex.sync()
        ex.sync()
        ms = (time.perf_counter() - t0) / a.steps * 1e3
        print(f"{name:<13} chunks {a.chunks:>2} {'sync every step' if per_step_sync else 'sync at the end '}: {ms:.3f} ms / step", flush=True)
ex.kernel_ms()
ex.run_async(a.steps)
ex.sync()
print("plain batch of steps: %.3f ms / step" % ex.kernel_ms()[0])
for _ in range(a.steps):
    ex.run_async(1)
ex.sync()
print("plain single-step launches: %.3f ms / step" % ex.kernel_ms()[0])
