#!/bin/bash
# usage (under gpurun): tools/gpu_variants.sh <tag> <ENV_NAME> "<v1 v2 ...>" [extra bench args]  -- one short bench per value
TAG=$1; NAME=$2; VALS=$3; shift 3
mkdir -p gpurun_out
for v in $VALS; do
  for st in warm cold; do
    env $NAME=$v python bench.py --no-extras --no-cpu-baseline --adam-state $st "$@" > gpurun_out/${TAG}_${NAME}_${v}_${st}.json 2> gpurun_out/${TAG}_${NAME}_${v}_${st}.err || tail -3 gpurun_out/${TAG}_${NAME}_${v}_${st}.err
    echo "== $NAME=$v $st"; python tools/bench_brief.py gpurun_out/${TAG}_${NAME}_${v}_${st}.json | head -2
  done
done
