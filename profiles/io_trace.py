#!/usr/bin/env python
"""Device timeline of one end-to-end step (mdr_phys_step_io_async) through MDR_PHYS_IO_TRACE=1: per world range, when
its copy-in landed, its kernel ended and its copy-out ended.  python profiles/io_trace.py [--chunks 6] [--no-in] [--no-out]"""
import argparse
import os
import sys

os.environ.setdefault("MDR_PHYS_IO_TRACE", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_ecs_madrona_b200 import phys, scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--chunks", type=int, default=6)
ap.add_argument("--no-in", action="store_true")
ap.add_argument("--no-out", action="store_true")
a = ap.parse_args()
sc = scenes.collisions64(16384, max_contacts=640)
ex = phys.PhysExecutor(sc)
ex.run(130)
ins = [] if a.no_in else [(c, ex.pinned(c)) for c in (phys.COL_EXT_FORCE, phys.COL_EXT_TORQUE)]
outs = [] if a.no_out else [(c, ex.pinned(c)) for c in (phys.COL_POSITION, phys.COL_ROTATION, phys.COL_VELOCITY)]
for _, arr in ins:
    arr[:] = 0
for _ in range(3):
    ex.step_io_async(ins, outs, a.chunks)
    ex.sync()
