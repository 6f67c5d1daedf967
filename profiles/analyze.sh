#!/bin/bash
# Usage: profiles/analyze.sh <tag> <kernel mangled substring>   (run in the build container)
TAG=$1; K=$2
mkdir -p /tmp/cub && cd /tmp/cub && rm -f *.cubin && cuobjdump -xelf all /root/repo/gpu_ecs_madrona_b200/libmadrona_b200.so > /dev/null
for f in *.cubin; do nvdisasm -gi -c $f; done > dis.txt 2>/dev/null
cd /root/repo
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_sass.csv 2>/dev/null
python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/${TAG}_raw.csv')))
hdr,units,vals=rows[0],rows[1],rows[2]
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__inst_executed.avg.per_cycle_elapsed','smsp__inst_executed.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','smsp__thread_inst_executed_per_inst_executed.ratio','launch__occupancy_limit','sass__inst_executed_local','launch__grid_size','launch__block_size','smsp__inst_executed_pipe_xu','smsp__inst_executed_pipe_fma.sum','smsp__inst_executed_pipe_alu.sum','smsp__inst_executed_pipe_fmaheavy','smsp__inst_executed_pipe_lsu.sum']
for i,h in enumerate(hdr):
    if any(h.startswith(w) for w in want) and 'pct' not in h.replace('pct_of_peak_sustained_active','') and 'per_second' not in h:
        print(f"{h:<70}{units[i]:<12}{vals[i]}")
PY
python profiles/attribute.py /tmp/cub/dis.txt gpurun_out/${TAG}_sass.csv "$K"
