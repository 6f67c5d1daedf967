#!/bin/bash
# ncu capture of the BATCHED ECS step (12 steps in one launch) of the three built-in worlds. Usage: profiles/run_ncu_ecs3.sh <tag>
# Launches of `bench_ecs.py --app X --steps 12 --warmup 30`: the warm-up batch, the timed batch (captured), 12 single steps.
TAG=${1:-r02f}
mkdir -p gpurun_out
for app in fvs fanout taskgraph; do
  CMD="python bench_ecs.py --app $app --steps 12 --warmup 30 --no-cpu"
  $CMD > gpurun_out/${TAG}_ecs_${app}_plain.json 2> gpurun_out/${TAG}_ecs_${app}_plain.err || { tail -3 gpurun_out/${TAG}_ecs_${app}_plain.err; continue; }
  ncu --set full --clock-control none -k regex:ecs.*StepKernel -s 1 -c 1 -f -o gpurun_out/${TAG}_ecs_${app} $CMD > gpurun_out/${TAG}_ecs_${app}_ncu.log 2>&1
  ncu -i gpurun_out/${TAG}_ecs_${app}.ncu-rep --page raw --csv > gpurun_out/${TAG}_ecs_${app}_raw.csv 2>/dev/null
done
python profiles/ncu_summary.py gpurun_out/${TAG}_ecs_fvs_raw.csv gpurun_out/${TAG}_ecs_fanout_raw.csv gpurun_out/${TAG}_ecs_taskgraph_raw.csv
