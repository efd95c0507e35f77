#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s -k "random_background or libtorch_shim" > gpurun_out/r02h_pytest.log 2>&1; tail -4 gpurun_out/r02h_pytest.log
python bench.py --mode render --config cfg3_4m_1080p_sh3 > gpurun_out/r02h_cfg3.json 2> gpurun_out/r02h_cfg3.err || tail -3 gpurun_out/r02h_cfg3.err
head -c 1200 gpurun_out/r02h_cfg3.json; echo
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; timeout 300 $TR bench.py --gpus 2 "$@" > gpurun_out/r02h_$name.json 2> gpurun_out/r02h_$name.err || tail -5 gpurun_out/r02h_$name.err; echo "== $name"; python tools/bench_brief.py gpurun_out/r02h_$name.json | head -2; grep -o '"replicas_identical": [a-z]*, "last_loss": [0-9.e-]*, "barrier_timeouts": [0-9a-z]*' gpurun_out/r02h_$name.json | head -2; grep -o '"exchange_ms": [0-9.]*' gpurun_out/r02h_$name.json; grep -o '"barrier_timeouts": [0-9a-z]*' gpurun_out/r02h_$name.json | tail -1; }
run peer_flags --steps 20 --warmup 5 --no-extras --no-cpu-baseline
GS_PEER_NCCL_BARRIER=1 run peer_nccl --steps 20 --warmup 5 --no-extras --no-cpu-baseline
run cfg4_pull --steps 3 --warmup 2 --batch-views 64
