#!/usr/bin/env python
"""Per-phase share of a block's time in the step kernel (clock64 between the block barriers of stepWorld;
needs a -DMDR_PHASE_CLOCKS experiment build: profiles/exp/build_exp.sh NAME -DMDR_PHASE_CLOCKS).

  python profiles/phases.py --lib libexp_NAME.so [--threads T --wpb B ...]
"""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NAMES = ["load", "leaf update + refit", "findPairs + levels", "integrate + boxes", "narrowphase pass", "hull-hull warps",
         "build level lists", "solve positions", "joints + setVelocities", "solve velocities", "contact count",
         "end leaf update", "store"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", required=True)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--wpb", type=int, default=0)
    ap.add_argument("--smem-contacts", type=int, default=0)
    ap.add_argument("--max-contacts", type=int, default=640)
    ap.add_argument("--worlds", type=int, default=16384)
    a = ap.parse_args()
    os.environ["MDR_LIB"] = a.lib
    from gpu_ecs_madrona_b200 import phys, scenes
    sc = scenes.collisions64(a.worlds, max_contacts=a.max_contacts)
    ex = phys.PhysExecutor(sc, block_threads=a.threads, worlds_per_block=a.wpb, smem_contacts=a.smem_contacts)
    ex.lib.mdr_phys_phase_clocks.argtypes = [C.c_void_p, C.c_void_p]
    buf = (C.c_ulonglong * 32)()
    info = ex.launch_info()

    def window(steps, label):
        ex.lib.mdr_phys_phase_clocks(ex.h, buf)
        ex.kernel_ms()
        ex.run_async(steps)
        ex.sync()
        ms, k = ex.kernel_ms()
        ex.lib.mdr_phys_phase_clocks(ex.h, buf)
        tot = sum(buf[:13])
        rounds = a.worlds / (info["blocks"] * info["worlds_per_block"])
        per = steps * info["blocks"] * rounds
        print(f"== {label}: {ms:.3f} ms/step, {a.worlds / ms / 1e3:.2f} M world-steps/s; "
              f"{tot / per:.0f} cycles per world-step of a block")
        for i, n in enumerate(NAMES):
            print(f"  {n:<26}{100.0 * buf[i] / tot:6.1f} %   {buf[i] / per:9.0f} cycles")
        for i, n in zip(range(16, 21), ("pair tests", "tail rows + slot store", "scan", "pair fill", "row order + levels")):
            print(f"    findPairs: {n:<22}{buf[i] / per:9.0f} cycles")

    ex.run(5)
    window(20, "falling (steps 5-25)")
    ex.run(85)
    window(20, "settled (steps 110-130)")
    ex.close()


if __name__ == "__main__":
    main()
