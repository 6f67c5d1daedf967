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


PREROLL = 120    # untimed steps before warm-up: every cube has landed and come to rest (contacts per world plateau at 64
                 # from step 100 on, ms per step within 1 % afterwards: profiles/r02i_step_profile.txt)


def cpu_baseline(steps, warmup, worlds_per_thread=256, kind=None, preroll=PREROLL, min_seconds=1.5):
    """Times the reference's CPU implementation of the path on this host's cores:
    the unmodified reference compiled into oracle/_ref when it travelled here,
    else the oracle restatement. A bounded sample of the same workload: the same
    per-world scene in the same regime (the same untimed pre-roll and warm-up), fewer
    worlds, and at least `min_seconds` of timed wall clock (the settled scene is
    stationary, so extra steps time the same work)."""
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
    o.step(preroll + warmup, cores)
    timed = 0
    t0 = time.perf_counter()
    while True:
        o.step(steps, cores)
        timed += steps
        dt = time.perf_counter() - t0
        if dt >= min_seconds:
            break
    o.close()
    return {"value": worlds * timed / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{worlds} worlds x {timed} steps after {preroll} pre-roll + {warmup} warm-up steps (settled regime, "
                      f"as the GPU arm), collisions-64, {cores} threads ({dt:.1f} s wall); worlds capped at 1024 by the "
                      "reference's 48 GB address-space reservation per world"}, dt / timed


def mean_pairs_contacts(args, gpu):
    """Mean candidate pairs / contacts per world-step over the TIMED step range, from an untimed
    replay of the first 256 worlds of the batch (same scene, same steps), sampled every step:
    the inputs of the algorithmic-byte figure of SURVEY 8(d)."""
    from gpu_ecs_madrona_b200 import phys, scenes
    sc = scenes.collisions64(256, max_contacts=args.max_contacts)
    ex = phys.PhysExecutor(sc, gpu_id=gpu)
    ex.run(args.preroll + args.warmup)
    ps, cs = [], []
    for _ in range(args.steps):
        ex.run(1)
        p, c = ex.stats()
        ps.append(p.mean())
        cs.append(c.mean())
    ex.close()
    return float(np.mean(ps)), float(np.mean(cs))


def timed_window(scene, gpu, skip, steps, lib=None, **shape):
    """Device-timed M world-steps/s of `steps` steps after `skip` untimed ones, fresh executor (sub-records)."""
    from gpu_ecs_madrona_b200 import phys
    ex = phys.PhysExecutor(scene, gpu_id=gpu, lib=lib, **shape)
    ex.run(skip)
    ex.kernel_ms()
    ex.run_async(steps)
    ex.sync()
    ms, n = ex.kernel_ms()
    ex.close()
    return {"value": scene.num_worlds * 1e3 / ms, "unit": UNIT, "ms_per_step": ms, "steps": n, "after_steps": skip}


def extra_records(args, sc, local, world, rank, shape):
    """Sub-records of the line (rank 0 fills them; every rank runs the configs[4]-sized leg at N > 1)."""
    import ctypes as C
    from gpu_ecs_madrona_b200 import phys, scenes
    out = {}
    if rank == 0:
        # the two regimes of the scene: the headline is the settled one whatever --steps / --warmup say
        out["regimes"] = {"falling": dict(timed_window(sc, local, 5, 20, **shape), note="steps 5-25: a third of the cubes rest on the plane"),
                          "landing": dict(timed_window(sc, local, 80, 20, **shape), note="steps 80-100: the top layer lands, the heaviest steps of the episode"),
                          "settled": dict(timed_window(sc, local, args.preroll, 20, **shape), note="after the pre-roll: the headline's regime")}
        fma = os.path.join(os.path.dirname(phys.LIB_PATH), "libmadrona_b200_fma.so")
        if os.path.exists(fma):
            rec = timed_window(sc, local, args.preroll + args.warmup, args.steps, lib=phys.bind(C.CDLL(fma)), **shape)
            rec["build"] = ("-fmad=true (FMA-contracted; pair lists / counts bit-exact, positions and rotations <= 1e-5 after one "
                            "step, velocities NOT: tests/test_fma_build.py) -- reported, not the headline")
            out["fma_build"] = rec
    return out


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    base, s_per_step = cpu_baseline(args.steps, args.warmup, preroll=args.preroll)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * s_per_step, "higher_is_better": True, "scaling": "weak",
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
    ex.run(args.preroll + args.warmup)
    ex.kernel_ms()
    sampler.wait_ready()
    barrier()
    sampler.mark()
    t0 = time.perf_counter()
    ex.run_async(args.steps)
    ex.sync()
    wall = time.perf_counter() - t0
    kms, launches = ex.kernel_ms()           # CUDA events on the executor's stream around the K steps
    gpu_launches = ex.last_launches()        # one call of K steps = one batch launch of the step kernel
    clocks = sampler.stop()
    assert launches == args.steps
    dev_ms = kms * launches
    P, C = mean_pairs_contacts(args, local)
    if dist is not None:
        t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())
    value = world * W * args.steps / (dev_ms * 1e-3)
    # the same K steps as K calls of one step each (one launch per step, the form a caller that looks at the state
    # after every step uses); reported beside the batch, not instead of it
    for _ in range(args.steps):
        ex.run_async(1)
    ex.sync()
    kms1, n1 = ex.kernel_ms()
    one_ms = kms1 * n1
    if dist is not None:
        t = torch.tensor([one_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        one_ms = float(t.item())
    one_step_per_call = {"value": world * W * n1 / (one_ms * 1e-3), "unit": UNIT, "ms_per_step": one_ms / n1,
                         "gpu_launches": n1}

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

    ex.run(args.preroll)
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
        es.launch(sharding.local_statistics(ex, args.preroll + args.warmup + e2e_steps,
                                            outs[0][1].view(np.float32)[:, :N_ROWS]))
        episode = {k: float(v) for k, v in sharding.EpisodeStats.totals(es.result()).items()}
    except Exception as e:      # statistics must never cost the bench line
        episode = {"error": repr(e)}

    ex.close()
    shape = dict(block_threads=args.threads, worlds_per_block=args.worlds_per_block, smem_contacts=args.smem_contacts)
    extras = {}
    if not args.no_extra:
        extras = extra_records(args, sc, local, world, rank, shape)
        if world > 1:
            # BASELINE.json configs[4]: collisions-64 x 1 048 576 worlds sharded over the N GPUs of the job
            # (every rank steps its 1 M / N worlds; settled regime; device-timed, max over ranks)
            Wbig = (1 << 20) // world
            big = scenes.collisions64(Wbig, world_offset=rank * Wbig, max_contacts=args.max_contacts)
            rec = timed_window(big, local, args.preroll, 10, **shape)
            t = torch.tensor([rec["ms_per_step"]], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if rank == 0:
                ms = float(t.item())
                extras["configs4_1M_worlds"] = {"value": (1 << 20) * 1e3 / ms, "unit": UNIT, "ms_per_step": ms, "steps": 10,
                                                "worlds_per_gpu": Wbig, "after_steps": args.preroll,
                                                "workload": "BASELINE.json configs[4]: 1 048 576 worlds over %d GPUs" % world}
        elif rank == 0:
            import bench_ecs
            ecs = {}
            for name in ("taskgraph", "fanout", "fvs"):       # BASELINE.json configs[0], [1], [3]
                try:
                    ecs[name] = bench_ecs.run_record(name, steps=100, warmup=10, no_cpu=args.no_cpu)
                except Exception as e:                          # a sub-record must never cost the headline
                    ecs[name] = {"error": repr(e)}
            extras["ecs_configs"] = ecs

    if rank == 0:
        peak, peak_src = measured_peak()
        A = algorithmic_bytes(P, C)
        achieved = A * W / (kms * 1e-3) / 1e9
        # DRAM bytes per launch cannot be measured without a profiler: the figure is the STATIC one of the last
        # `ncu --set full` capture of this kernel in this regime (profiles/traffic.json names capture and build)
        traffic, traffic_src = None, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            try:
                tj = json.load(open(tpath))
                # per STEP: the timed region is one launch of `steps` steps (gpu_launches), the capture one of 12
                traffic, traffic_src = tj.get("dram_bytes_per_step"), "static: " + tj.get("source", "profiles/traffic.json")
            except Exception:
                traffic = None
        base, _ = cpu_baseline(args.steps, args.warmup, preroll=args.preroll) if not args.no_cpu else ({}, 0)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"collisions-64: {W} worlds per GPU x (64 dynamic unit cubes + 1 static "
                                   "plane), 4 substeps, dt 1/60 (BASELINE.json configs[2])",
                       "worlds_per_gpu": W, "bodies_per_world": N_ROWS, "substeps": SUBSTEPS,
                       "mean_pairs_per_world": P, "mean_contacts_per_world": C,
                       "regime": f"settled: {args.preroll} untimed pre-roll steps, then {args.warmup} warm-up, then the {args.steps} "
                                 "timed steps (contacts per world have plateaued; every --steps / --warmup pair times this regime)",
                       "l2": "state tables (%.0f MB per GPU) exceed the 126 MB L2; no flush needed"
                             % (W * 68 * 290 / 1e6),
                       "launch": info, "wall_ms_per_step": 1e3 * wall / args.steps,
                       "parallelism": f"worlds sharded over {world} GPU(s), no per-step communication",
                       "episode_stats_all_gather": episode},
            "mean_contacts_per_world": C, "mean_pairs_per_world": P,      # the regime the value was measured in
            "gpu_launches": gpu_launches, "steps_per_launch": args.steps // max(gpu_launches, 1),
            "one_step_per_call": one_step_per_call,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps, "finite": finite,
                    "pcie_gb_per_s_per_rank": (h2d + d2h) * e2e_steps / e2e_s / 1e9,
                    "columns": "in: ExternalForce, ExternalTorque; out: Position, Rotation, Velocity (whole column tiles)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes_per_world_step": A, "kernel": "physStepKernel",
                         "kernel_ms": kms,
                         "per": "step: achieved = A x worlds / device ms per step, traffic = DRAM bytes per step; the timed "
                                "launch holds steps_per_launch steps, so per launch both scale by that number"},
            "cpu_baseline": base,
            "clocks": clocks,
        }
        line.update(extras)
        print(json.dumps(line))
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
    ap.add_argument("--io-chunks", type=int, default=14,
                    help="world ranges the end-to-end step is pipelined over (1: copy, step, copy in sequence; "
                         "14 = one range per resident wave of 16384 worlds: the shortest exposed copy-out)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--preroll", type=int, default=PREROLL,
                    help="untimed steps before the warm-up, so that every --steps / --warmup pair times the settled regime")
    ap.add_argument("--no-extra", action="store_true",
                    help="skip the sub-records (regimes, FMA build, ECS configs, configs[4] size): the headline only")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
