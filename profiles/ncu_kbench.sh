#!/bin/bash
# Usage (under gpurun, 1 GPU): [ENV=...] profiles/ncu_kbench.sh TAG [kbench args]
# One `ncu --set full` capture of the step kernel in the settled regime (launch 116 of kbench's 130),
# after the same command has run plain. Writes gpurun_out/TAG_{plain.txt,prof.ncu-rep,raw.csv}.
TAG=$1; shift
mkdir -p gpurun_out
python profiles/kbench.py "$@" --tag $TAG > gpurun_out/${TAG}_plain.txt 2>&1 || { cat gpurun_out/${TAG}_plain.txt; exit 1; }
cat gpurun_out/${TAG}_plain.txt
ncu --set full --clock-control none --import-source on -k regex:physStepKernel -s ${SKIP:-115} -c 1 \
    -f -o gpurun_out/${TAG}_prof python profiles/kbench.py "$@" --tag $TAG > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
python profiles/ncu_summary.py gpurun_out/${TAG}_raw.csv
