#!/usr/bin/env python
"""ms per step of collisions-64 in 10-step windows from the start of the episode (where does the scene settle?)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_ecs_madrona_b200 import phys, scenes
sc = scenes.collisions64(16384)
ex = phys.PhysExecutor(sc)
for w in range(30):
    ex.kernel_ms(); ex.run_async(10); ex.sync()
    ms, n = ex.kernel_ms()
    p, c = ex.stats()
    print(f"steps {10*w:3d}-{10*w+10:3d}: {ms:.3f} ms/step  {16384/ms/1e3:.2f} M/s  pairs {p.mean():.1f} contacts {c.mean():.1f}")
