#!/bin/bash
# usage (under gpurun): tools/gpu_env_sweep.sh <tag> "<ENV=val ENV2=val2>" ...  -- one short default-config bench per setting
TAG=$1; shift
mkdir -p gpurun_out
i=0
for setting in "base" "$@"; do
  i=$((i+1))
  if [ "$setting" = "base" ]; then E=""; else E="$setting"; fi
  env $E python bench.py --no-extras --no-cpu-baseline > gpurun_out/${TAG}_$i.json 2> gpurun_out/${TAG}_$i.err || tail -3 gpurun_out/${TAG}_$i.err
  echo "== [$setting]"; python tools/bench_brief.py gpurun_out/${TAG}_$i.json | head -2 | cut -c1-300
done
