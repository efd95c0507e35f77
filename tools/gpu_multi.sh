#!/bin/bash
# usage (under gpurun --gpus N): tools/gpu_multi.sh <tag> <N>  -- the data-parallel modes at N GPUs, short runs
TAG=$1; N=$2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; $TR bench.py --gpus $N "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err || tail -5 gpurun_out/${TAG}_$name.err; echo "== $name"; python tools/bench_brief.py gpurun_out/${TAG}_$name.json | head -3; grep -o '"replicas_identical": [a-z]*, "last_loss": [0-9.e-]*' gpurun_out/${TAG}_$name.json | head -2; grep -o '"exchange_ms": [0-9.]*' gpurun_out/${TAG}_$name.json; }
run peer --steps 10 --warmup 3 --no-extras --no-cpu-baseline
run allreduce --steps 10 --warmup 3 --no-extras --no-cpu-baseline --dp-mode allreduce
run batch8_pull --steps 3 --warmup 2 --batch-views 8 --config cfg2_1p6m_1246x825_sh3
run batch8_allreduce --steps 3 --warmup 2 --batch-views 8 --config cfg2_1p6m_1246x825_sh3 --dp-mode allreduce
run cfg4_pull --steps 3 --warmup 2 --batch-views 64
run cfg4_allreduce --steps 3 --warmup 2 --batch-views 64 --dp-mode allreduce
