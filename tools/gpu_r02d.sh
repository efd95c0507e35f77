#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s -k "checkpoint or native or capacity or adam" > gpurun_out/r02d_pytest.log 2>&1; tail -8 gpurun_out/r02d_pytest.log
tools/gpu_variants.sh r02d GS_FUSED_VARIANT "0 4 5 6 7"
export GS_FUSED_NO_L2_PREFETCH=1
tools/gpu_variants.sh r02d_nopf GS_FUSED_VARIANT "0 4 6"
