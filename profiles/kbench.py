#!/usr/bin/env python
"""Kernel-only timing of the collisions-64 step for launch-shape / build sweeps.

  python profiles/kbench.py [--lib NAME.so] [--threads T --wpb B --smem-contacts C --max-contacts M]
                            [--worlds W] [--tag TEXT]

Prints one line: device-timed M world-steps/s of three step windows of ONE run,
steps 5-25 (falling), 10-110 (BASELINE.md protocol: 100 after 10) and 110-130 (settled).
CUDA events on the executor's stream (mdr_phys_kernel_ms). Not a bench line: bench.py is.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default="")
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--wpb", type=int, default=0)
    ap.add_argument("--smem-contacts", type=int, default=0)
    ap.add_argument("--max-contacts", type=int, default=640)
    ap.add_argument("--worlds", type=int, default=16384)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    if a.lib:
        os.environ["MDR_LIB"] = a.lib
    from gpu_ecs_madrona_b200 import phys, scenes
    sc = scenes.collisions64(a.worlds, max_contacts=a.max_contacts)
    ex = phys.PhysExecutor(sc, block_threads=a.threads, worlds_per_block=a.wpb, smem_contacts=a.smem_contacts)
    info = ex.launch_info()

    def timed(n):
        ex.kernel_ms()
        ex.run_async(n)
        ex.sync()
        ms, k = ex.kernel_ms()
        assert k == n
        return ms * k

    ex.run(5)
    t_a = timed(5)            # steps 5-10
    t_b = timed(15)           # steps 10-25
    t_c = timed(85)           # steps 25-110
    t_d = timed(20)           # steps 110-130
    W = a.worlds
    falling = W * 20 / ((t_a + t_b) * 1e-3) / 1e6
    proto = W * 100 / ((t_b + t_c) * 1e-3) / 1e6
    settled = W * 20 / (t_d * 1e-3) / 1e6
    print(f"{a.tag or a.lib or 'default':<28} falling {falling:6.2f}  protocol {proto:6.2f}  settled {settled:6.2f}  M world-steps/s  "
          f"| {info['blocks']}x{info['threads']} thr, {info['blocks_per_sm']} blk/SM, {info['threads_per_world']} thr/world, "
          f"{info['worlds_per_block']} worlds/blk, smem {info['smem_bytes']}", flush=True)
    ex.close()


if __name__ == "__main__":
    main()
