#!/bin/bash
# Usage: profiles/exp/build_exp.sh NAME [extra nvcc flags...]   -> gpu_ecs_madrona_b200/libexp_NAME.so
# physics.cu + the experiment instantiation unit only (no ECS / sort): for kernel experiments with
# profiles/kbench.py --lib libexp_NAME.so. VARIANTS="256,2 256,3" lists (max threads,min blocks) variants;
# FMAD=1 in the environment builds with -fmad=true.
set -e
NAME=$1; shift
cd "$(dirname "$0")/../.."
V=${VARIANTS:-"256,2"}
LIST=$(for v in $V; do printf 'MDR_STEP_VARIANT(%s), ' "$v"; done)
echo "#define MDR_EXP_LIST $LIST" > profiles/exp/exp_list.h
FM=false; [ "${FMAD:-0}" = 1 ] && FM=true
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=$FM -prec-div=true -prec-sqrt=true \
  -ftz=false -Xcompiler -fPIC,-ffp-contract=off -shared --threads 0 -Xptxas -v "$@" \
  gpu_ecs_madrona_b200/csrc/physics.cu profiles/exp/phys_inst_exp.cu \
  -o gpu_ecs_madrona_b200/libexp_$NAME.so 2>&1 | grep -A2 "physStepKernel" | grep -E "Used|spill" | sed 's/^ *//'
