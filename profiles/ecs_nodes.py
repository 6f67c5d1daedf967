#!/usr/bin/env python
"""Per-node share of a warp's walk of fantasy_vs (clock64 around every node; needs a -DMDR_ECS_NODE_CLOCKS build of the
library:  nvcc ... -DMDR_ECS_NODE_CLOCKS csrc/*.cu -o libecsclk.so).   python profiles/ecs_nodes.py --lib libecsclk.so"""
import argparse, ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ap = argparse.ArgumentParser(); ap.add_argument("--lib", required=True); ap.add_argument("--worlds", type=int, default=8192)
a = ap.parse_args()
os.environ["MDR_LIB"] = a.lib
from gpu_ecs_madrona_b200 import ecs_apps as A
from gpu_ecs_madrona_b200.ecs import EcsExecutor
ex = EcsExecutor(A.APP_FVS, a.worlds)
ex.run(30)
buf = (C.c_ulonglong * 16)()
ex.lib.mdr_ecs_node_clocks.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
ex.lib.mdr_ecs_node_clocks(ex.h, buf, 16)
ex.step_ms(); ex.run(20)
ms, n = ex.step_ms()
ex.lib.mdr_ecs_node_clocks(ex.h, buf, 16)
names = ["action", "caster (+ nested blasts)", "archer", "mark", "sort tracker", "destroy", "respawn", "clear tracker", "tick"]
tot = sum(buf[:9])
print(f"{ms:.4f} ms / step; {tot / 20 / a.worlds:.0f} cycles per world-step of a warp")
for i, nm in enumerate(names):
    print(f"  {nm:<26}{100.0 * buf[i] / tot:6.1f} %  {buf[i] / 20 / a.worlds:9.0f} cycles")
