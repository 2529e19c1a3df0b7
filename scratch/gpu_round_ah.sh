#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -2 gpurun_out/pytest_gpu.log
SRT_UPLOAD_LOG=1 SRT_BENCH_DEBUG=1 timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu --workload soup > gpurun_out/bench_ah_soup.json 2> gpurun_out/log_ah_soup.txt; grep -v "^wave" gpurun_out/log_ah_soup.txt | tail -5
python -c "import json; d=json.load(open('gpurun_out/bench_ah_soup.json')); print('soup', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],2), round(d['e2e']['ms_per_step'],1))"
