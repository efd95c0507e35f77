#!/bin/bash
# usage (under gpurun): tools/gpu_check.sh <tag> [pytest -k expression]  -- GPU parity tests, then the default bench
TAG=$1; K=${2:-}
mkdir -p gpurun_out
if [ -n "$K" ]; then python -m pytest tests -m gpu -x -q -k "$K" > gpurun_out/${TAG}_pytest.log 2>&1; else python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; fi
tail -4 gpurun_out/${TAG}_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.log 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
python tools/bench_brief.py gpurun_out/${TAG}_bench.log
