#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu "$@" > gpurun_out/e_$name.json 2> gpurun_out/e_$name.err; python -c "import json; d=json.load(open('gpurun_out/e_$name.json')); print('$name', round(d['value'],1), round(d['ms_per_step'],2), round(d['e2e']['value'],1), d['gpu_launches'])" >> gpurun_out/e_summary.txt 2>&1; }
rm -f gpurun_out/e_summary.txt
run s32m --wave-slots 33554432
run s64m --wave-slots 67108864
run s134m --wave-slots 134217728
for r in 1 2 4 8 12 16 24; do run s16m_r$r --wave-slots 16777216 --refill $r; done
run s16m_b32 --wave-slots 16777216 --block 32
run s16m_b128 --wave-slots 16777216 --block 128
run s16m_ref --wave-slots 16777216 --traversal reference
cat gpurun_out/e_summary.txt
