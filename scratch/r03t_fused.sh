#!/bin/bash
# round 3t: SRT_OPT_FUSED_PRIMARY -- parity first, then the option's two values on Rubik (whole frame, an eighth) and the stand-in
cd "$(dirname "$0")/.."
timeout 500 python -m pytest tests/test_fused_primary_gpu.py tests/test_finder_build_gpu.py tests/test_checked_build_gpu.py -x -q -m gpu 2>&1 | tail -15
for fz in 0 1; do
  for so in 1 8; do
    timeout 300 python bench.py --workload rubik --steps 10 --warmup 3 --no-cpu --stripe-of $so --fused-primary $fz > gpurun_out/r03t_f${fz}_of$so.json 2>gpurun_out/r03t_f${fz}_of$so.err
    python -c "import json;d=json.load(open('gpurun_out/r03t_f${fz}_of$so.json'));print('fused $fz stripe-of $so:', round(d['ms_per_step'],3),'ms', d['value'])"
  done
done
