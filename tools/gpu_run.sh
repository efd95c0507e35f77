#!/bin/bash
# usage (under gpurun): tools/gpu_run.sh <tag> [pytest -k expr | "none"] [ncu kernel regex | "none"]
# GPU tests (all failures shown, no -x), the default bench, then the ncu launch list and one --set full capture.
TAG=$1; K=${2:-}; PAT=${3:-none}
mkdir -p gpurun_out
if [ "$K" != "none" ]; then
  if [ -n "$K" ]; then python -m pytest tests -m gpu -q -s -k "$K" > gpurun_out/${TAG}_pytest.log 2>&1; else python -m pytest tests -m gpu -q -s > gpurun_out/${TAG}_pytest.log 2>&1; fi
  tail -15 gpurun_out/${TAG}_pytest.log
fi
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
python tools/bench_brief.py gpurun_out/${TAG}_bench.json
[ "$PAT" = "none" ] && exit 0
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/${TAG}_plain.log 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -20 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 160 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"$PAT" -s 6 -c 2 \
    -f -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -3 gpurun_out/${TAG}_ncu_full.log
ls -la gpurun_out/ | tail -12
