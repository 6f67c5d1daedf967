#!/bin/bash
# ncu capture of the ECS megakernel on fantasy_vs with per-source-line attribution. Usage: profiles/run_ncu_ecs2.sh <tag>
TAG=${1:-r02p}
CMD="python bench_ecs.py --app fvs --steps 12 --warmup 30 --no-cpu"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:ecsFusedStepKernel -s 32 -c 1 \
    -f -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page source --csv --print-source cuda > gpurun_out/${TAG}_src.csv 2>/dev/null
ls -la gpurun_out | grep ${TAG}
