#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
run() {  # tag, extra args
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu $2 > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err
  python -c "import json; d=json.load(open('gpurun_out/bench_$1.json')); print('$1', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1))" || tail -5 gpurun_out/bench_$1.err
}
run af_rubik ""
run af_air "--workload airplane-standin"
SRT_WAVE_LOG=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu > /dev/null 2> gpurun_out/wave_log_af.txt; grep "^wave" gpurun_out/wave_log_af.txt | tail -10
