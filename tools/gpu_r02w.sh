#!/bin/bash
T=${1:-r02w}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
python bench.py --no-cpu-baseline --no-extras > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || tail -5 gpurun_out/${T}_bench.err
python tools/bench_brief.py gpurun_out/${T}_bench.json | head -2 | cut -c1-400
python bench.py --no-cpu-baseline --no-extras --variant stress > gpurun_out/${T}_stress.json 2> gpurun_out/${T}_stress.err || tail -5 gpurun_out/${T}_stress.err
python tools/bench_brief.py gpurun_out/${T}_stress.json | head -2 | cut -c1-400
