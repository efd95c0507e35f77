#!/bin/bash
T=${1:-r02n}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || tail -5 gpurun_out/${T}_bench.err
python tools/bench_brief.py gpurun_out/${T}_bench.json | cut -c1-1200
grep -o '"host_enqueue_ms_per_step": [0-9.]*' gpurun_out/${T}_bench.json | head -3
