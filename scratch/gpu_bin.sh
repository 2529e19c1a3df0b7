#!/bin/bash
mkdir -p gpurun_out
for b in 0 1000000; do
SRT_WAVE_LOG=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --bin-rays $b > gpurun_out/bin_$b.json 2> gpurun_out/bin_$b.err
python -c "import json; d=json.load(open('gpurun_out/bin_$b.json')); print('bin $b', round(d['value'],1), round(d['ms_per_step'],2))"
grep "^wave [0-4]:" gpurun_out/bin_$b.err | tail -5
done
