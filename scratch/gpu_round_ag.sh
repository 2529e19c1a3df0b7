#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -2 gpurun_out/pytest_gpu.log
SRT_UPLOAD_LOG=1 SRT_BENCH_DEBUG=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --workload airplane-standin > gpurun_out/bench_ag_air.json 2> gpurun_out/wave_log_air_ag.txt; tail -6 gpurun_out/wave_log_air_ag.txt
python -c "import json; d=json.load(open('gpurun_out/bench_ag_air.json')); print('air', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2))"
