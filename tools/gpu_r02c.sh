#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/r02c_pytest.log 2>&1; tail -12 gpurun_out/r02c_pytest.log
tools/gpu_variants.sh r02c GS_FUSED_VARIANT "0 2 3"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
GS_FUSED_VARIANT=2 $CMD > gpurun_out/r02c_plain.log 2>&1 && GS_FUSED_VARIANT=2 ncu --set full --clock-control none --import-source on -k regex:preprocess_backward_adam -s 6 -c 1 -f -o gpurun_out/r02c_v2 $CMD > gpurun_out/r02c_ncu.log 2>&1
GS_FUSED_VARIANT=0 ncu --set full --clock-control none --import-source on -k regex:preprocess_backward_adam -s 6 -c 1 -f -o gpurun_out/r02c_v0 $CMD >> gpurun_out/r02c_ncu.log 2>&1
tail -2 gpurun_out/r02c_ncu.log
