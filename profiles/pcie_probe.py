#!/usr/bin/env python
"""Host <-> device copy rates of the box (pinned memory, CUDA events): one direction alone, both at once, and a
device-to-host copy split over 1 / 2 / 4 streams. Answers whether the end-to-end floor of bench.py (57.9 MB out per
step) is the link or one copy engine.  python profiles/pcie_probe.py"""
import torch

MB = 58
n = MB * 1024 * 1024
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
dev2 = torch.empty(n // 2, dtype=torch.uint8, device="cuda")
host = torch.empty(n, dtype=torch.uint8).pin_memory()
host2 = torch.empty(n // 2, dtype=torch.uint8).pin_memory()


def timed(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def d2h_split(k):
    streams = [torch.cuda.Stream() for _ in range(k)]
    chunk = n // k

    def fn():
        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)
        done = []
        for i, s in enumerate(streams):
            s.wait_event(ev)
            with torch.cuda.stream(s):
                host[i * chunk:(i + 1) * chunk].copy_(dev[i * chunk:(i + 1) * chunk], non_blocking=True)
                e = torch.cuda.Event()
                e.record(s)
                done.append(e)
        for e in done:
            cur.wait_event(e)
    return fn


ms = timed(lambda: host.copy_(dev, non_blocking=True))
print(f"D2H alone, {MB} MB: {ms:.3f} ms = {n / ms / 1e6:.1f} GB/s")
ms = timed(lambda: dev.copy_(host, non_blocking=True))
print(f"H2D alone, {MB} MB: {ms:.3f} ms = {n / ms / 1e6:.1f} GB/s")
side = torch.cuda.Stream()


def both():
    cur = torch.cuda.current_stream()
    ev = torch.cuda.Event()
    ev.record(cur)
    side.wait_event(ev)
    with torch.cuda.stream(side):
        dev2.copy_(host2, non_blocking=True)
        e = torch.cuda.Event()
        e.record(side)
    host.copy_(dev, non_blocking=True)
    cur.wait_event(e)


ms = timed(both)
print(f"D2H {MB} MB + H2D {MB // 2} MB at once: {ms:.3f} ms = {1.5 * n / ms / 1e6:.1f} GB/s in total")
for k in (1, 2, 4):
    ms = timed(d2h_split(k))
    print(f"D2H {MB} MB over {k} stream(s): {ms:.3f} ms = {n / ms / 1e6:.1f} GB/s")
