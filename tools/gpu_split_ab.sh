#!/bin/bash
# usage (under gpurun --gpus N): tools/gpu_split_ab.sh <tag> <N> [test]  -- split modes (copy engines / SM kernel / off),
# short lines with the phase timings
TAG=$1; N=$2
mkdir -p gpurun_out
if [ "${3:-}" = "test" ]; then
  timeout 600 python -m pytest tests/test_gpu_split_phase.py -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
fi
TR="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for v in ${MODES:-ce off}; do
  GS_PEER_SPLIT=$v GS_PEER_NO_SPLIT=$([ $v = off ] && echo 1 || echo 0) $TR bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${TAG}_$v.json 2> gpurun_out/${TAG}_$v.err || tail -8 gpurun_out/${TAG}_$v.err
  echo "== split mode $v"; python tools/bench_brief.py gpurun_out/${TAG}_$v.json | head -2 | cut -c1-400; grep -o '"peer_phases_ms": {.*}}' gpurun_out/${TAG}_$v.json | cut -c1-640; grep -o '"replicas_identical": [a-z]*, "last_loss": [0-9.e-]*, "barrier_timeouts": [0-9a-z]*' gpurun_out/${TAG}_$v.json
done
