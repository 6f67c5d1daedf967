#!/bin/bash
# Round-2 profiling pass of the step kernel (run under gpurun, 1 GPU). Usage: profiles/run_ncu_r02.sh <tag>
# plain run first, then the launch list of the same command, then one `--set full` capture of a settled launch.
set -u
TAG=${1:-r02m}
CMD="python bench.py --steps 12 --warmup 5 --no-cpu --no-extra"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
cut -c1-300 gpurun_out/${TAG}_plain.json
# (since batches: run(preroll + warm-up) is ONE launch of 125 steps, the timed region ONE launch of 12 steps, then 12
#  single-step launches of the one_step_per_call record, then the world ranges of the end-to-end steps)
ncu --metrics gpu__time_duration.sum --clock-control none -s 0 -c 120 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:physStepKernel -s 1 -c 1 \
    -f -o gpurun_out/${TAG}_prof $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page source --csv --print-source sass > gpurun_out/${TAG}_sass.csv 2>/dev/null
python profiles/ncu_summary.py gpurun_out/${TAG}_raw.csv
ls -la gpurun_out | grep ${TAG}
