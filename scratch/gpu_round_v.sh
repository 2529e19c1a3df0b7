#!/bin/bash
# round 1v: tail kernel occupancy variants, out-of-line sincos
mkdir -p gpurun_out
run() {  # tag, extra args
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu $2 > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err
  python -c "import json; d=json.load(open('gpurun_out/bench_$1.json')); print('$1', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1))" || tail -5 gpurun_out/bench_$1.err
}
run v_base ""
run v_base_air "--workload airplane-standin"
cp simple-ray-tracer_b200/libsrt_b200.so /tmp/keep.so
for v in tail13 tail16 sincos; do cp scratch/libsrt_$v.so simple-ray-tracer_b200/libsrt_b200.so; run v_$v ""; run v_${v}_air "--workload airplane-standin"; done
cp scratch/libsrt_sincos.so simple-ray-tracer_b200/libsrt_b200.so
timeout 900 python -m pytest tests -m gpu -x -q -k "rubik_reference or airplane or two_models" > gpurun_out/pytest_sincos.log 2>&1; tail -2 gpurun_out/pytest_sincos.log
cp /tmp/keep.so simple-ray-tracer_b200/libsrt_b200.so
