#!/bin/bash
# round 3aa: find kernel at 12 blocks per SM (75-80 registers, no spills) against 16 (64 registers, 24-64 B of spills)
cd "$(dirname "$0")/.."
for lib in "" scratch/variants/libsrt_find12.so; do
  name=$([ -z "$lib" ] && echo b16 || echo b12)
  for wl in rubik airplane-standin; do
    SRT_LIB=$([ -z "$lib" ] || echo $PWD/$lib) timeout 100 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu > gpurun_out/r03aa_${name}_${wl}.json 2>gpurun_out/r03aa_${name}_${wl}.err
    python -c "import json;d=json.load(open('gpurun_out/r03aa_${name}_${wl}.json'));print('$name $wl:', round(d['ms_per_step'],3),'ms', round(d['value'],1), 'trace ms', round(d['roofline']['kernel_ms_total'],2))"
  done
done
