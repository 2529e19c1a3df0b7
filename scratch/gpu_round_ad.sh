#!/bin/bash
mkdir -p gpurun_out
run() {  # tag, extra args
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu $2 > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err
  python -c "import json; d=json.load(open('gpurun_out/bench_$1.json')); print('$1', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1))" || tail -5 gpurun_out/bench_$1.err
}
run ad_base ""
run ad_base_air "--workload airplane-standin"
cp simple-ray-tracer_b200/libsrt_b200.so /tmp/keep.so
for v in logic3x256 logic6x128 logic8x128; do cp scratch/libsrt_$v.so simple-ray-tracer_b200/libsrt_b200.so; run ad_$v ""; run ad_${v}_air "--workload airplane-standin"; done
cp /tmp/keep.so simple-ray-tracer_b200/libsrt_b200.so
