#!/bin/bash
T=${1:-r02v}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_split_phase.py tests/test_zz_no_alloc.py tests/test_gpu_parity.py -m gpu -x -q -k "split or colour or tiny or static or small_scene or baseline_configs or stage" > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || tail -5 gpurun_out/${T}_bench.err
python tools/bench_brief.py gpurun_out/${T}_bench.json | head -3 | cut -c1-700
grep -o '"clocks": {[^}]*}' gpurun_out/${T}_bench.json | head -3
