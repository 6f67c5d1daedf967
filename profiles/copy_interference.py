#!/usr/bin/env python
"""Does a device-to-host copy running beside the step kernel slow the kernel down? A batch of steps timed alone, with a
side stream copying 58 MB device -> pinned host in a loop, and with a host -> device loop.  python profiles/copy_interference.py"""
import os
import sys
import threading

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_ecs_madrona_b200 import phys, scenes  # noqa: E402

sc = scenes.collisions64(16384, max_contacts=640)
ex = phys.PhysExecutor(sc)
ex.run(130)
n = 58 * 1024 * 1024
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
host = torch.empty(n, dtype=torch.uint8).pin_memory()
side = torch.cuda.Stream()


def batch(tag, copies):
    ex.kernel_ms()
    ex.run_async(40)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 30
    if copies:
        with torch.cuda.stream(side):
            a.record(side)
            for _ in range(reps):        # ~32 ms of copies (alone) beside ~70 ms of steps
                if copies == "d2h":
                    host.copy_(dev, non_blocking=True)
                else:
                    dev.copy_(host, non_blocking=True)
            b.record(side)
    ex.sync()
    side.synchronize()
    rate = f", copies at {n * reps / a.elapsed_time(b) / 1e6:.1f} GB/s" if copies else ""
    print(f"{tag:<28} {ex.kernel_ms()[0]:.3f} ms / step{rate}", flush=True)


with torch.cuda.stream(side):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(side)
    for _ in range(30):
        host.copy_(dev, non_blocking=True)
    e1.record(side)
side.synchronize()
print(f"device -> host copies alone: {n * 30 / e0.elapsed_time(e1) / 1e6:.1f} GB/s")
batch("alone", None)
batch("beside device -> host copies", "d2h")
batch("beside host -> device copies", "h2d")
batch("alone", None)
