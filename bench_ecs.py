#!/usr/bin/env python
"""Secondary benchmark: the ECS / task-graph configs of BASELINE.json (configs[0], [1], [3]) on
the CUDA executor, with the reference's CPU code timed beside it. Not the driver's bench line
(that is bench.py, collisions-64); the numbers feed DESIGN.md and profiles/.

  python bench_ecs.py [--app fvs|fanout|taskgraph|all] [--worlds N] [--steps K] [--warmup W]

A step = one MWCudaExecutor::run(): export copy-in, the task-graph megakernel over every world,
export offsets + copy-out. Device-timed with CUDA events on the executor's stream.
Algorithmic bytes per world-step follow SURVEY 8(d): per node, (component bytes read + written
per row) x rows.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from gpu_ecs_madrona_b200 import ecs_apps as A  # noqa: E402

APPS = {"taskgraph": (A.APP_TASKGRAPH, 1024), "fanout": (A.APP_FANOUT, 4096), "fvs": (A.APP_FVS, 8192)}


def algorithmic_bytes(app, rows):
    """Graph-faithful traffic per world-step (bytes), node by node; rows = mean row counts."""
    if app == A.APP_TASKGRAPH:      # drift R+W Position, clamp R+W Position, export 28 B
        n = rows[0] + rows[1]
        return 2 * 24 * n + 28 * rows[1]
    if app == A.APP_FANOUT:         # move 24, box 12+16 R + 24 W, overlap N*(8+24) + N^2/2*24 + P*20, sort, consume
        n, p = rows[0], rows[1]
        return 24 * n + 52 * n + 32 * n + 12 * n * n + 20 * p + 4 * 2 * 20 * p + p * (16 + 2 * 12 + 2 * 8) + 4 * n
    if app == A.APP_FVS:            # action 8+12+4 RW, caster, archer, mark: per row
        nd, nk = rows[0], rows[1]
        n = nd + nk
        return n * (8 + 2 * 12 + 2 * 4) + nd * (8 + 2 * 4 + 2 * 4) + nk * (8 + 2 * 4 + 2 * 4) + n * (8 + 4) + 4 * nd + 12 * nk
    return 0


def cpu_baseline(app, warmup, worlds_per_thread, min_seconds=1.2):
    """The reference's own state layer + task graph (oracle/_ref/libmadrona_ecs_ref.so) on this host's cores,
    timed for at least `min_seconds` of wall clock (steps added in batches until then)."""
    from oracle import eorc
    kind = "reference" if eorc.available("reference") else "port"
    cores = len(os.sched_getaffinity(0))
    worlds = worlds_per_thread * cores
    o = eorc.make_app_oracle(kind, app, worlds)
    o.step(warmup, cores)
    steps, batch = 0, 20
    t0 = time.perf_counter()
    while True:
        o.step(batch, cores)
        steps += batch
        dt = time.perf_counter() - t0
        if dt >= min_seconds:
            break
        batch = min(4 * batch, 2000)
    o.close()
    return {"value": worlds * steps / dt, "unit": "world-steps/s", "cores": cores, "kind": kind,
            "sample": f"{worlds} worlds x {steps} steps after {warmup} warm-up, {cores} threads ({dt:.2f} s wall)"}


def run_record(name, worlds=0, steps=100, warmup=10, no_cpu=False):
    """One ECS config on the CUDA executor: device-timed value, end-to-end value (every exported column read
    back to host memory after every step), roofline on the graph-faithful bytes, CPU baseline. Returns a dict."""
    from gpu_ecs_madrona_b200.ecs import EcsExecutor
    app, dflt = APPS[name]
    W = worlds or dflt
    ex = EcsExecutor(app, W)
    ex.run(warmup)
    ex.step_ms()
    ex.run_async(steps)          # ONE call of `steps` steps: a batch, walked inside one launch (mdr_ecs_last_launches)
    ex.sync()
    ms, n = ex.step_ms()
    launches = ex.last_launches()
    # the same number of steps as `steps` calls of one step each (one graph launch per step), for comparison
    for _ in range(steps):
        ex.run_async(1)
    ex.sync()
    ms1, n1 = ex.step_ms()
    launches1 = n1 * ex.launches_per_step()
    rows = ex.row_counts().mean(axis=0)
    Abytes = algorithmic_bytes(app, rows)
    peak = 6463.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    # end to end: a fresh executor over the same step range, one run() + host copies of the exports per step
    ex.close()
    ex = EcsExecutor(app, W)
    slots = U_NUM_EXPORTS[app]
    ex.run(warmup)
    d2h = 0
    bufs = [ex.pinned_export_buffer(s) for s in range(slots)]     # page-locked, allocated once
    t0 = time.perf_counter()
    for _ in range(steps):
        d2h = sum(ex.step_read(list(range(slots)), bufs))       # one step + every exported column in host memory
    e2e_s = time.perf_counter() - t0
    line = {"metric": f"world-steps/sec ({A.APP_NAMES[app]}, device-timed)", "value": W * n / (ms * n * 1e-3),
            "unit": "world-steps/s", "n_gpus": 1, "steps": n, "warmup": warmup, "ms_per_step": ms,
            "config": {"workload": f"{A.APP_NAMES[app]}: {W} worlds, params {ex.params}", "mean_rows": rows.tolist()},
            "gpu_launches": launches, "steps_per_call": n,
            "one_step_per_call": {"value": W / (ms1 * 1e-3), "ms_per_step": ms1, "gpu_launches": launches1},
            "e2e": {"value": W * steps / e2e_s, "unit": "world-steps/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": int(d2h), "steps": steps},
            "roofline": {"bound": "hbm", "achieved": Abytes * W / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": Abytes * W / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_world_step": Abytes,
                         "kernel": "ecsStepKernel (+export kernels)", "traffic": None}}
    if not no_cpu:
        line["cpu_baseline"] = cpu_baseline(app, min(warmup, 5), 8)
    ex.close()
    return line


U_NUM_EXPORTS = {A.APP_TASKGRAPH: 2, A.APP_FANOUT: 1, A.APP_FVS: 2}


def run(name, worlds, steps, warmup, no_cpu):
    print(json.dumps(run_record(name, worlds, steps, warmup, no_cpu)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--app", default="all")
    ap.add_argument("--worlds", type=int, default=0)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    for name in (APPS if args.app == "all" else [args.app]):
        run(name, args.worlds, args.steps, args.warmup, args.no_cpu)


if __name__ == "__main__":
    main()
