#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 --workload airplane-standin > gpurun_out/bench_airplane_r01f.json 2> gpurun_out/bench_airplane.err; echo "airplane rc=$?"
timeout 1500 python bench.py --steps 3 --warmup 3 --workload soup --soup-tris 10000000 --cpu-frames 1 > gpurun_out/bench_soup10m_r01f.json 2> gpurun_out/bench_soup.err; echo "soup rc=$?"
for f in airplane soup10m; do python -c "import json; d=json.load(open('gpurun_out/bench_${f}_r01f.json')); print('$f', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['roofline']['frac'], d['cpu_baseline'])"; done
