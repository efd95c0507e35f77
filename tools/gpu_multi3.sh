#!/bin/bash
# usage (under gpurun --gpus N): tools/gpu_multi3.sh <tag> <N>  -- the driver's own multi-GPU command only
TAG=$1; N=$2
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${TAG}_default.json 2> gpurun_out/${TAG}_default.err || tail -8 gpurun_out/${TAG}_default.err
python tools/bench_brief.py gpurun_out/${TAG}_default.json | cut -c1-1600
grep -o '"replicas_identical": [a-z]*, "last_loss": [0-9.e-]*, "barrier_timeouts": [0-9a-z]*' gpurun_out/${TAG}_default.json | head -2
