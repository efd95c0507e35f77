#!/bin/bash
# usage: tools/profile_kernel.sh <tag> <kernel-regex> [skip] [count]   -- one --set full capture of matching kernels
set -u
TAG=$1; PAT=$2; SKIP=${3:-8}; CNT=${4:-2}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -20 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$PAT" -s $SKIP -c $CNT \
    -f -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_ncu.log
