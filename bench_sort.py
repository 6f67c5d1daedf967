#!/usr/bin/env python
"""Bench of the global archetype sort / compaction (mdr_sort_*), not the
driver's line (bench.py is): the collisions-64 rigid-body archetype stored as ONE shared table
(Entity, WorldID, 12 phys::Cols columns: 192 B per row), rows in creation order with 20 % created
late and 10 % destroyed, sorted by WorldID into a second set of column buffers.

Algorithmic bytes per row: histogram 4 + passes x 16 (key + index, read + write)
+ gather 2 x 192 + 4 (index) = 424 B at 2 passes. Device time: CUDA events on the operator's stream.
Prints one JSON line per table size."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PHYS_COLS = [8, 4, 12, 16, 12, 24, 4, 4, 28, 28, 24, 12, 12, 4]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, nargs="*", default=[16384, 131072])
    ap.add_argument("--rows-per-world", type=int, default=65)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--lib", default="", help="another build of csrc/table_sort.cu (profiles/exp sweeps)")
    args = ap.parse_args()
    import torch
    from gpu_ecs_madrona_b200 import table_sort
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        src = "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        peak, src = 6650.0, "fallback of B200_PROFILING.md"
    for W in args.worlds:
        cols = table_sort.churned_table(W, args.rows_per_world, PHYS_COLS, seed=1)
        n = len(cols[1])
        d_in = [torch.from_numpy(np.ascontiguousarray(c)).cuda() for c in cols]
        d_out = [torch.empty_like(c) for c in d_in]
        d_map = torch.full((n,), -1, dtype=torch.int32, device="cuda")
        import ctypes
        s = table_sort.TableSort(n, PHYS_COLS, 1, W, lib=ctypes.CDLL(args.lib) if args.lib else None)
        for _ in range(args.warmup):
            rows = s.run(d_in, d_out, n, d_map)
        ms, launches = [], 0
        for _ in range(args.steps):
            rows = s.run(d_in, d_out, n, d_map)       # the input is never modified: same work every step
            t, launches = s.last_ms()
            ms.append(t)
        row_bytes = sum(PHYS_COLS)
        A = n * (4 + s.passes * 16 + 2 * row_bytes + 4 + 8 + 4)      # + remap: entity read, map write
        med = float(np.median(ms))
        line = {"metric": "rows sorted/sec (global archetype sort by WorldID, device-timed)",
                "value": n / (med * 1e-3), "unit": "rows/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": med, "dtype": "u32",
                "config": {"workload": f"collisions-64 archetype as one shared table: {W} worlds x "
                                       f"{args.rows_per_world} rows, {row_bytes} B/row, 20 % late rows, 10 % destroyed",
                           "rows": n, "rows_kept": rows, "passes": s.passes,
                           "l2": f"table {n * row_bytes / 1e6:.0f} MB x 2 buffers"},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": A / (med * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": A / (med * 1e-3) / 1e9 / peak, "peak_source": src,
                             "algorithmic_bytes_per_row": A / n, "kernel": "tableGatherKernel + sortOnesweepKernel"}}
        if not args.no_cpu and n <= 4_000_000:
            from oracle import sort_oracle
            t0 = time.perf_counter()
            sort_oracle.sort_table(cols, 1, W, np.full(n, -1, np.int32))
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / dt, "unit": "rows/s", "cores": 1, "kind": "port",
                                    "sample": f"the same table once, numpy restatement ({dt:.2f} s)"}
        print(json.dumps(line), flush=True)
        s.close()
        del d_in, d_out, d_map
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
