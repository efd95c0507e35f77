#!/bin/bash
# usage (under gpurun --gpus N): tools/gpu_split.sh <tag> <N> [notest]  -- split-phase step: unit tests on one GPU, then
# the driver's multi-GPU command with the split (default) and without (GS_PEER_NO_SPLIT=1), short lines
TAG=$1; N=$2
mkdir -p gpurun_out
if [ "${3:-}" != "notest" ]; then
  timeout 600 python -m pytest tests/test_gpu_split_phase.py tests/test_gpu_parity.py -k "split or colour or virtual_ranks or small_scene" -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -5 gpurun_out/${TAG}_pytest.log
fi
TR="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; $TR bench.py --gpus $N "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err || tail -8 gpurun_out/${TAG}_$name.err; echo "== $name"; python tools/bench_brief.py gpurun_out/${TAG}_$name.json | cut -c1-900; grep -o '"replicas_identical": [a-z]*, "last_loss": [0-9.e-]*, "barrier_timeouts": [0-9a-z]*' gpurun_out/${TAG}_$name.json | head -2; }
run split --steps 20 --warmup 3 --no-cpu-baseline --sustain-seconds 1
GS_PEER_NO_SPLIT=1 run nosplit --steps 20 --warmup 3 --no-cpu-baseline --no-extras
