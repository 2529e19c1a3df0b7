#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
: > gpurun_out/sweep.txt
for cfg in "$@"; do
  IFS=, read r a l w <<< "$cfg"
  v=$(( r | (a<<8) | (l<<16) | (w<<24) ))
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --refill $v > gpurun_out/sw.json 2> gpurun_out/sw.err
  python -c "import json; d=json.load(open('gpurun_out/sw.json')); print('cfg $cfg', round(d['value'],1), round(d['ms_per_step'],2))" | tee -a gpurun_out/sweep.txt
done
