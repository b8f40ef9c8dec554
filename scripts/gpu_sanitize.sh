#!/bin/bash
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 120 python scripts/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 500 compute-sanitizer --tool $TOOL --print-limit 20 python scripts/sanitize_target.py > gpurun_out/sanitize_$TOOL.log 2>&1; echo "sanitizer rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize target ok|Invalid|Race|hazard" gpurun_out/sanitize_$TOOL.log | head -20
