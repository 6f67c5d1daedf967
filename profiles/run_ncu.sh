#!/bin/bash
# Profiling pass of the step kernel on the GPU box (run under gpurun, 1 GPU).
# Usage: profiles/run_ncu.sh <tag> [extra bench args]
# Writes gpurun_out/<tag>_{plain.log,launches.csv,prof.ncu-rep,raw.csv}.
set -u
TAG=${1:-r01}; shift || true
CMD="python bench.py --steps 12 --warmup 60 --no-cpu $*"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
tail -1 gpurun_out/${TAG}_plain.log | cut -c1-400
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 40 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:physStepKernel -s 62 -c 1 \
    -f -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ls -la gpurun_out | grep ${TAG}
