#!/bin/bash
# round 3g: the stand-in's tail threshold per rank of 8 (its reference tree has leaves of up to 165 triangles: the tail's walk is slow)
cd "$(dirname "$0")/.."
for tp in 1048576 262144 65536 16384 0; do
  for so in 1 8; do
    timeout 600 python bench.py --workload airplane-standin --steps 6 --warmup 3 --no-cpu --tail-paths $tp --stripe-of $so > gpurun_out/r03g_standin_tail${tp}_of$so.json 2>/dev/null
    python -c "import json;d=json.load(open('gpurun_out/r03g_standin_tail${tp}_of$so.json'));print('stand-in tail $tp stripe-of $so:', round(d['ms_per_step'],3),'ms; launches/step', d['gpu_launches']/d['steps'])"
  done
done
