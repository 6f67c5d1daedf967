#!/usr/bin/env python
"""Key metrics of one or more `ncu --page raw --csv` files, side by side.  python profiles/ncu_summary.py a_raw.csv [b_raw.csv ...]"""
import csv
import sys

WANT = [
    ("ms", "gpu__time_duration.sum"),
    ("warp inst (M)", "smsp__inst_executed.sum"),
    ("threads / inst", "smsp__thread_inst_executed_per_inst_executed.ratio"),
    ("% issue active", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("% warps active", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("registers", "launch__registers_per_thread"),
    ("grid", "launch__grid_size"), ("block", "launch__block_size"),
    ("dyn smem / block", "launch__shared_mem_per_block_dynamic"),
    ("MB read", "dram__bytes_read.sum"), ("MB written", "dram__bytes_write.sum"),
    ("local ld inst", "sass__inst_executed_local_loads"), ("local st inst", "sass__inst_executed_local_stores"),
    ("smem bank conflicts", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
    ("l1 hit %", "l1tex__t_sector_hit_rate.pct"),
    ("% fma pipe", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    ("% alu pipe", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    ("% lsu pipe", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
    ("% lsu wavefronts", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
]
STALLS = ["barrier", "wait", "long_scoreboard", "short_scoreboard", "not_selected", "no_instruction",
          "branch_resolving", "math_pipe_throttle", "mio_throttle", "lg_throttle", "dispatch_stall",
          "membar", "sleeping", "tex_throttle", "drain", "misc"]


def load(path):
    rows = list(csv.reader(open(path)))
    hdr, vals = rows[0], rows[2]
    return {h: v for h, v in zip(hdr, vals)}


def main():
    files = sys.argv[1:]
    data = [load(f) for f in files]
    print("| metric | " + " | ".join(f.split("/")[-1].replace("_raw.csv", "") for f in files) + " |")
    print("|---|" + "---|" * len(files))

    def row(label, key, scale=1.0):
        cells = []
        for d in data:
            v = d.get(key, "")
            try:
                x = float(v.replace(",", "")) * scale
                cells.append(f"{x:.4g}")
            except ValueError:
                cells.append(v or "-")
        print(f"| {label} | " + " | ".join(cells) + " |")

    for label, key in WANT:
        row(label, key, 1e-6 if label.startswith("warp inst") else 1.0)
    for s in STALLS:
        row(f"stall {s} / issue", f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio")


if __name__ == "__main__":
    main()
