#!/bin/bash
T=${1:-r02o}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_raw_path.py tests/test_gpu_split_phase.py -m gpu -x -q -k "adam or train or fused or split or peer or virtual" > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || tail -5 gpurun_out/${T}_bench.err
python tools/bench_brief.py gpurun_out/${T}_bench.json | cut -c1-1500 | grep -v "^   north\|^   roofline\|^   stress\|^   best"
python bench.py --no-cpu-baseline --no-extras --no-fused > gpurun_out/${T}_nofused.json 2> gpurun_out/${T}_nofused.err || tail -5 gpurun_out/${T}_nofused.err
python tools/bench_brief.py gpurun_out/${T}_nofused.json | head -2
