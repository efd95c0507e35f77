#!/bin/bash
T=${1:-r02p}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_raw_path.py tests/test_gpu_split_phase.py -m gpu -x -q -k "adam or train or fused or split or peer or virtual or tiny" > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
CMD="python bench.py --no-cpu-baseline --no-extras --no-fused --steps 4 --warmup 3"
python bench.py --no-cpu-baseline --no-extras --no-fused > gpurun_out/${T}_nofused.json 2> gpurun_out/${T}_nofused.err || tail -5 gpurun_out/${T}_nofused.err
python tools/bench_brief.py gpurun_out/${T}_nofused.json | head -2
ncu --set full --clock-control none --import-source on -k regex:'adam_stats' -s 12 -c 2 -f -o gpurun_out/${T}_adam_stats $CMD > gpurun_out/${T}_ncu.log 2>&1; tail -2 gpurun_out/${T}_ncu.log
