#!/usr/bin/env python
"""Secondary benchmark: batched BVH::traceRay (the lidar query) on the
collisions-64 worlds, with the reference's own traceRay timed on the host cores beside it.
Not the driver's bench line (bench.py).

  python bench_rays.py [--worlds 16384] [--rays 64] [--steps 50] [--warmup 5] [--settle 40]

A step = one mdr_phys_trace_rays_device launch over every world (rays resident in HBM);
device-timed with CUDA events on the executor's stream. Algorithmic bytes per launch:
28 B in + 28 B out per ray (origin, direction, t_max; t, normal, entity, row = 28) and 80 B
per leaf read once per world.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def lidar_rays(W, R, pos, rng):
    """R rays per world fanned around a sensor 1.5 above the ground plane near the grid's centre."""
    centre = pos[:, 1:].mean(axis=1)
    o = np.repeat(centre[:, None, :], R, axis=1).astype(np.float32)
    o[:, :, 2] = 1.5
    o[:, :, :2] += rng.uniform(-4, 4, (W, 1, 2)).astype(np.float32)
    ang = (np.arange(R, dtype=np.float32) / R * 2 * np.pi)[None, :]
    d = np.stack([np.cos(ang), np.sin(ang), np.full_like(ang, -0.35)], axis=-1).astype(np.float32)
    return o, np.ascontiguousarray(np.broadcast_to(d, (W, R, 3)))


def cpu_baseline(sc_fn, settle, R, seconds=10.0):
    from oracle import orc
    kind = "reference" if orc.available("reference") else "port"
    cores = len(os.sched_getaffinity(0))
    W = min(64, 4 * cores)
    sc = sc_fn(W)
    o = orc.Oracle(sc, kind)
    o.step(settle, cores)
    rng = np.random.default_rng(1)
    og, dr = lidar_rays(W, R, o.read_f32(orc.COL_POSITION), rng)
    import threading
    done = [0] * cores

    def work(tid):
        t_end = time.perf_counter() + seconds
        n = 0
        while time.perf_counter() < t_end:
            for w in range(tid, W, cores):
                o.trace_rays(w, og[w], dr[w], np.float32(np.inf))      # ctypes releases the GIL
                n += R
        done[tid] = n
    t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(i,)) for i in range(cores)]
    [t.start() for t in th]
    [t.join() for t in th]
    dt = time.perf_counter() - t0
    return {"value": sum(done) / dt, "unit": "rays/s", "cores": cores, "kind": kind,
            "sample": f"{W} worlds x {R} rays repeated for {dt:.1f} s on {cores} threads"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=16384)
    ap.add_argument("--rays", type=int, default=64)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--settle", type=int, default=40, help="physics steps before the rays are cast")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    from gpu_ecs_madrona_b200 import phys, scenes
    from bench import measured_peak, ClockSampler
    W, R = args.worlds, args.rays
    sc = scenes.collisions64(W)
    ex = phys.PhysExecutor(sc)
    ex.run(args.settle)
    rng = np.random.default_rng(1)
    og, dr = lidar_rays(W, R, ex.state()[0], rng)
    d_o, d_d = torch.from_numpy(og).cuda(), torch.from_numpy(dr).cuda()
    out_t = torch.empty(W, R, device="cuda")
    out_n = torch.empty(W, R, 3, device="cuda")
    out_e = torch.empty(W, R, 2, dtype=torch.int32, device="cuda")
    out_r = torch.empty(W, R, dtype=torch.int32, device="cuda")
    stream = torch.cuda.ExternalStream(ex.lib.mdr_phys_stream(ex.h))
    torch.cuda.synchronize()

    def launch():
        rc = ex.lib.mdr_phys_trace_rays_device(ex.h, R, d_o.data_ptr(), d_d.data_ptr(), None, out_t.data_ptr(),
                                               out_n.data_ptr(), out_e.data_ptr(), out_r.data_ptr())
        assert rc == 0
    for _ in range(args.warmup):
        launch()
    ex.sync()
    sampler = ClockSampler(0)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(args.steps):
            launch()
        ev1.record(stream)
    ex.sync()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1) / args.steps
    hit = float((out_r >= 0).float().mean().item())
    # end to end: host rays in, host results out through the blocking host-pointer entry
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        ex.trace_rays(og, dr)
    e2e = W * R * reps / (time.perf_counter() - t0)
    peak, peak_src = measured_peak()
    A = W * R * 56 + W * 65 * 80
    line = {"metric": "rays/sec (BVH::traceRay, collisions-64 worlds, device-timed)", "value": W * R / (ms * 1e-3),
            "unit": "rays/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{W} worlds x {R} rays after {args.settle} physics steps, 65 leaves per world",
                       "hit_fraction": hit, "l2": "rays + outputs %.0f MB per launch" % (W * R * 56 / 1e6)},
            "gpu_launches": args.steps, "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": A / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": A / (ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src},
            "e2e": {"value": e2e, "unit": "rays/s", "h2d_bytes_per_step": W * R * 24, "d2h_bytes_per_step": W * R * 28}}
    if not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(scenes.collisions64, args.settle, R)
    print(json.dumps(line))


if __name__ == "__main__":
    main()
