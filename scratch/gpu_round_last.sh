#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -2 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/last_bench.json 2> gpurun_out/last_bench.err
python -c "import json; d=json.load(open('gpurun_out/last_bench.json')); print('default', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['cpu_baseline']['value'], d['clocks'])" || tail -5 gpurun_out/last_bench.err
SRT_UPLOAD_LOG=1 timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu --workload soup 2>&1 >/dev/null | grep upload_scene | tail -1
