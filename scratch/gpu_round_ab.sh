#!/bin/bash
mkdir -p gpurun_out
run() {  # tag, extra args
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu $2 > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err
  python -c "import json; d=json.load(open('gpurun_out/bench_$1.json')); print('$1', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1))" || tail -5 gpurun_out/bench_$1.err
}
for r in 4 8 12 16 20 24; do run ab_refill$r "--refill $r"; done
for r in 8 16 24; do run ab_air_refill$r "--refill $r --workload airplane-standin"; done
