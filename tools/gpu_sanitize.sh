#!/bin/bash
# usage (under gpurun): tools/gpu_sanitize.sh <tool>   (memcheck | racecheck | synccheck | initcheck; ONE tool per call)
TOOL=$1
mkdir -p gpurun_out
python tools/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
tail -1 gpurun_out/sanitize_plain.log
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python tools/sanitize_target.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize target ok" gpurun_out/sanitize_$TOOL.log; grep -c "=========     at" gpurun_out/sanitize_$TOOL.log; grep -E "========= (Invalid|Race|Error|Uninit|Barrier|Warning)" gpurun_out/sanitize_$TOOL.log | sort | uniq -c | head -20
