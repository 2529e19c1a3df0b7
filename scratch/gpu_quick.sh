#!/bin/bash
# parity suite + default bench (+ optional extra bench args via $1..)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
python -c "import json; d=json.load(open('gpurun_out/bench_quick.json')); print('BENCH', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'])"
SRT_WAVE_LOG=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu > /dev/null 2> gpurun_out/wave_log.txt; head -12 gpurun_out/wave_log.txt
