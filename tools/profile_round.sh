#!/bin/bash
# Profile pass on the GPU box (run under gpurun): plain run first, then the ncu launch list of the
# same command, then one --set full capture of every kernel of one steady-state step.
# usage: tools/profile_round.sh <tag> [launches]   (second argument: launch list only, no --set full pass)
set -u
TAG=${1:-rXX}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -20 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 120 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
[ "${2:-}" = "launches" ] && exit 0
ncu --set full --clock-control none --import-source on \
    -k regex:'preprocess|duplicate|identify|blend|adam|l1_loss|order_tiles|radix|scan|ssim' -s 136 -c 36 \
    -f -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
ls -la gpurun_out/
