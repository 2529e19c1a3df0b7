#!/bin/bash
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_bvh_build_gpu.py tests/test_parity_gpu.py -m gpu -x -q -k "not 300000 and not soup_depth8 and not inside_view" > gpurun_out/sanitizer_memcheck.log 2>&1; echo "rc=$?" >> gpurun_out/sanitizer_memcheck.log
tail -15 gpurun_out/sanitizer_memcheck.log
grep -c "Invalid\|ERROR SUMMARY" gpurun_out/sanitizer_memcheck.log
