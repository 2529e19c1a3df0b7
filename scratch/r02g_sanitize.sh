#!/bin/bash
# compute-sanitizer over scratch/sanitize_driver.py: memcheck, racecheck, initcheck; logs under gpurun_out/
cd "$(dirname "$0")/.."
python tests/kernel_driver.py > gpurun_out/r02g_driver_plain.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/r02g_driver_plain.log
for tool in memcheck racecheck initcheck; do
  t0=$(date +%s)
  timeout 900 compute-sanitizer --tool $tool --print-limit 40 python tests/kernel_driver.py > gpurun_out/r02g_sanitize_$tool.log 2>&1
  echo "$tool rc=$? $(( $(date +%s) - t0 )) s"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|ALL OK|SOME CASES" gpurun_out/r02g_sanitize_$tool.log | tail -3
done
