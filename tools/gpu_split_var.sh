#!/bin/bash
# usage (under gpurun --gpus N): tools/gpu_split_var.sh <tag> <N> "<grid values>"  -- SM split-phase step at several caps of
# the SH-phase grid (GS_PEER_ADAM_GRID), short lines with the phase timings
TAG=$1; N=$2; VALS=$3
mkdir -p gpurun_out
TR="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for v in $VALS; do
  GS_PEER_SPLIT=sm GS_PEER_ADAM_GRID=$v $TR bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${TAG}_g$v.json 2> gpurun_out/${TAG}_g$v.err || tail -8 gpurun_out/${TAG}_g$v.err
  echo "== grid $v"; python tools/bench_brief.py gpurun_out/${TAG}_g$v.json | head -2 | cut -c1-400; grep -o '"peer_phases_ms": {.*}}' gpurun_out/${TAG}_g$v.json | cut -c1-640
done
