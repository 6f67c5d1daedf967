#!/usr/bin/env python
"""Benchmark of the batched world step (BASELINE.json: aggregate world-steps/s,
collisions-64, 16 384 worlds per GPU).

  python bench.py [--gpus N] [--steps K] [--warmup W]        our CUDA path
  python bench.py --impl reference ...                       the reference's CPU path

A step = one pass of the whole physics task graph over every world of the batch.
Prints ONE JSON line (rank 0). See DESIGN.md "Measurement" for the definitions.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "aggregate world-steps/sec (collisions-64, device-timed)"
UNIT = "world-steps/s"
N_ROWS, SUBSTEPS = 65, 4


def algorithmic_bytes(pairs, contacts):
    """A = 480 N + 32 P + S (272 N + 104 P + 840 C) bytes per world-step."""
    return 480 * N_ROWS + 32 * pairs + SUBSTEPS * (272 * N_ROWS + 104 * pairs + 840 * contacts)


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
        except Exception:
            pass
    return 6650.0, "fallback of B200_PROFILING.md (6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def wait_ready(self, timeout=8.0):
        """nvidia-smi needs 0.1-1 s to come up (longer with 8 ranks starting one each): the sampler is
        started during the warm-up and the timed region begins once the first row has arrived."""
        t0 = time.perf_counter()
        while self.proc and not self.rows and time.perf_counter() - t0 < timeout:
            time.sleep(0.02)

    def mark(self):
        """Start of the timed region: earlier rows (warm-up, also under load) are dropped if later ones exist."""
        self.first = len(self.rows)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.t.join(timeout=2)
        first = getattr(self, "first", 0)
        if len(self.rows) > first:
            self.rows = self.rows[first:]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def cpu_baseline(steps, warmup, worlds_per_thread=256, kind=None):
    """Times the reference's CPU implementation of the path on this host's cores:
    the unmodified reference compiled into oracle/_ref when it travelled here,
    else the oracle restatement. A bounded sample of the same workload: the same
    per-world scene and the same step range, fewer worlds."""
    from gpu_ecs_madrona_b200 import scenes
    from oracle import orc
    if kind is None:
        kind = "reference" if orc.available("reference") else "port"
    if kind == "port" and not orc.available("port"):
        orc.build("port")
    cores = len(os.sched_getaffinity(0))
    # The reference reserves 48 GB of virtual address space per StateManager
    # (src/common/virtual.cpp:79), one per world: stay well under the 128 TB user space.
    worlds = min(worlds_per_thread * cores, 1024)
    worlds = max(cores, worlds // cores * cores)
    sc = scenes.collisions64(worlds)
    o = orc.Oracle(sc, kind)
    o.step(warmup, cores)
    t0 = time.perf_counter()
    o.step(steps, cores)
    dt = time.perf_counter() - t0
    o.close()
    return {"value": worlds * steps / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{worlds} worlds x {steps} steps after {warmup} warm-up steps (the same step range), "
                      f"collisions-64, {cores} threads ({dt:.1f} s wall); worlds capped at 1024 by the "
                      "reference's 48 GB address-space reservation per world"}, dt


def mean_pairs_contacts(args, gpu):
    """Mean candidate pairs / contacts per world-step over the TIMED step range, from an untimed
    replay of the first 256 worlds of the batch (same scene, same steps), sampled every step:
    the inputs of the algorithmic-byte figure of SURVEY 8(d)."""
    from gpu_ecs_madrona_b200 import phys, scenes
    sc = scenes.collisions64(256, max_contacts=args.max_contacts)
    ex = phys.PhysExecutor(sc, gpu_id=gpu)
    ex.run(args.warmup)
    ps, cs = [], []
    for _ in range(args.steps):
        ex.run(1)
        p, c = ex.stats()
        ps.append(p.mean())
        cs.append(c.mean())
    ex.close()
    return float(np.mean(ps)), float(np.mean(cs))


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    base, dt = cpu_baseline(args.steps, args.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "collisions-64 (64 dynamic unit cubes + 1 static plane per world, "
                                   "4 substeps, dt 1/60), bounded sample: " + base["sample"]},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def run_ours(args):
    rank, world, local = dist_env()
    import torch
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from gpu_ecs_madrona_b200 import phys, scenes
    W = args.worlds_per_gpu
    sc = scenes.collisions64(W, world_offset=rank * W, max_contacts=args.max_contacts)
    ex = phys.PhysExecutor(sc, gpu_id=local, block_threads=args.threads,
                           worlds_per_block=args.worlds_per_block, smem_contacts=args.smem_contacts)
    info = ex.launch_info()

    def barrier():
        if dist is not None:
            dist.barrier()
        ex.sync()

    # ---- device-resident throughput -------------------------------------
    sampler = ClockSampler(local)
    sampler.start()
    ex.run(args.warmup)
    ex.kernel_ms()
    sampler.wait_ready()
    barrier()
    sampler.mark()
    t0 = time.perf_counter()
    ex.run_async(args.steps)
    ex.sync()
    wall = time.perf_counter() - t0
    kms, launches = ex.kernel_ms()           # CUDA events on the executor's stream around the K launches
    clocks = sampler.stop()
    assert launches == args.steps
    dev_ms = kms * launches
    P, C = mean_pairs_contacts(args, local)
    if dist is not None:
        t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())
    value = world * W * args.steps / (dev_ms * 1e-3)

    # ---- end to end through the public API with host buffers ------------
    # A fresh executor on the same scene, so that the end-to-end figure covers the SAME step range
    # as `value` (warm-up steps, then K timed steps): the cost of a step depends on how far the cubes
    # have fallen (falling 1.4 ms, settled 2.8 ms), and a region tacked on after the device-timed one
    # would time a different workload.
    ex.close()
    ex = phys.PhysExecutor(sc, gpu_id=local, block_threads=args.threads,
                           worlds_per_block=args.worlds_per_block, smem_contacts=args.smem_contacts)
    # per step: actions in (ExternalForce, ExternalTorque) from pinned host memory,
    # one step, observations out (Position, Rotation, Velocity) to pinned host memory.
    ins = [(c, ex.pinned(c)) for c in (phys.COL_EXT_FORCE, phys.COL_EXT_TORQUE)]
    outs = [(c, ex.pinned(c)) for c in (phys.COL_POSITION, phys.COL_ROTATION, phys.COL_VELOCITY)]
    for _, a in ins:
        a[:] = 0
    h2d = sum(a.nbytes for _, a in ins)
    d2h = sum(a.nbytes for _, a in outs)

    def e2e_step():
        if args.io_chunks > 1:
            ex.step_io_async(ins, outs, args.io_chunks)     # copies overlap the step, range by range
        else:
            for c, a in ins:
                ex.write_raw_async(c, a)
            ex.run_async(1)
            for c, a in outs:
                ex.read_raw_async(c, a)
        ex.sync()

    for _ in range(args.warmup):
        e2e_step()
    barrier()
    e2e_steps = args.steps
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    e2e_s = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = world * W * e2e_steps / e2e_s
    finite = bool(np.isfinite(outs[0][1].view(np.float32)[:, :N_ROWS]).all())

    # ---- episode statistics: the path's one (optional) collective, off the step -------------
    # every rank's small statistics vector all-gathered over NCCL; not timed.
    episode = None
    try:
        from gpu_ecs_madrona_b200 import sharding
        es = sharding.EpisodeStats(every=args.steps, device=f"cuda:{local}")
        es.launch(sharding.local_statistics(ex, args.warmup + e2e_steps,
                                            outs[0][1].view(np.float32)[:, :N_ROWS]))
        episode = {k: float(v) for k, v in sharding.EpisodeStats.totals(es.result()).items()}
    except Exception as e:      # statistics must never cost the bench line
        episode = {"error": repr(e)}

    if rank == 0:
        peak, peak_src = measured_peak()
        A = algorithmic_bytes(P, C)
        achieved = A * W / (kms * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
            except Exception:
                traffic = None
        base, _ = cpu_baseline(args.steps, args.warmup) if not args.no_cpu else ({}, 0)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"collisions-64: {W} worlds per GPU x (64 dynamic unit cubes + 1 static "
                                   "plane), 4 substeps, dt 1/60 (BASELINE.json configs[2])",
                       "worlds_per_gpu": W, "bodies_per_world": N_ROWS, "substeps": SUBSTEPS,
                       "mean_pairs_per_world": P, "mean_contacts_per_world": C,
                       "l2": "state tables (%.0f MB per GPU) exceed the 126 MB L2; no flush needed"
                             % (W * ex.row_stride() * 290 / 1e6),
                       "launch": info, "wall_ms_per_step": 1e3 * wall / args.steps,
                       "parallelism": f"worlds sharded over {world} GPU(s), no per-step communication",
                       "episode_stats_all_gather": episode},
            "gpu_launches": args.steps,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps, "finite": finite},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_world_step": A, "kernel": "physStepKernel",
                         "kernel_ms": kms},
            "cpu_baseline": base,
            "clocks": clocks,
        }
        print(json.dumps(line))
    ex.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--worlds-per-gpu", type=int, default=16384)
    ap.add_argument("--threads", type=int, default=0, help="threads per world (0: library default)")
    ap.add_argument("--worlds-per-block", type=int, default=0)
    ap.add_argument("--smem-contacts", type=int, default=0, help="contact slots in shared memory (0: default)")
    ap.add_argument("--max-contacts", type=int, default=640, help="candidate-pair capacity per world (SURVEY 8d: 640)")
    ap.add_argument("--io-chunks", type=int, default=4,
                    help="world ranges the end-to-end step is pipelined over (1: copy, step, copy in sequence)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
