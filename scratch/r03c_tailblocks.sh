#!/bin/bash
# round 3c: blocks per SM of the tail kernel while the next chunk's waves run beside it
cd "$(dirname "$0")/.."
for tb in 0 3 5 7 9; do
  for so in 1 8; do
    SRT_TAIL_BLOCKS_PER_SM=$tb timeout 600 python bench.py --workload rubik --steps 10 --warmup 3 --no-cpu --stripe-of $so > gpurun_out/r03c_tb${tb}_of$so.json 2>/dev/null
    python -c "import json;d=json.load(open('gpurun_out/r03c_tb${tb}_of$so.json'));print('tail blocks/SM $tb stripe-of $so:', round(d['ms_per_step'],3),'ms')"
  done
  SRT_TAIL_BLOCKS_PER_SM=$tb timeout 600 python bench.py --workload airplane-standin --steps 10 --warmup 3 --no-cpu > gpurun_out/r03c_tb${tb}_standin.json 2>/dev/null
  python -c "import json;d=json.load(open('gpurun_out/r03c_tb${tb}_standin.json'));print('tail blocks/SM $tb stand-in:', round(d['ms_per_step'],3),'ms')"
done
