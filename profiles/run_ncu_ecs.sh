#!/bin/bash
# Profiling pass of the ECS task-graph megakernel (fantasy_vs, 8192 worlds). Usage: profiles/run_ncu_ecs.sh <tag>
set -u
TAG=${1:-r01e}
CMD="python bench_ecs.py --app fvs --steps 12 --warmup 30 --no-cpu"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
tail -1 gpurun_out/${TAG}_plain.log | cut -c1-600
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ecsStepKernel -s 32 -c 1 \
    -f -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ls -la gpurun_out | grep ${TAG}
