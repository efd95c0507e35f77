#!/bin/bash
# usage (under gpurun --gpus N): tools/gpu_multi2.sh <tag> <N>  -- the driver's own multi-GPU command (default flags, all
# extras: sustained + north_star cfg4 batch), then the NCCL all-reduce mode for the loss comparison
TAG=$1; N=$2
mkdir -p gpurun_out
TR="timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; $TR bench.py --gpus $N "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err || tail -8 gpurun_out/${TAG}_$name.err; echo "== $name"; python tools/bench_brief.py gpurun_out/${TAG}_$name.json | cut -c1-1500; grep -o '"replicas_identical": [a-z]*, "last_loss": [0-9.e-]*, "barrier_timeouts": [0-9a-z]*' gpurun_out/${TAG}_$name.json | head -2; }
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
run default --steps 20 --warmup 3
run allreduce --steps 20 --warmup 3 --no-extras --no-cpu-baseline --dp-mode allreduce
