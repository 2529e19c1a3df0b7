#!/bin/bash
# round 1n: hit finder -- parity suite, then finder on/off on the three workloads
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for f in 0 1; do
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --finder $f > gpurun_out/bench_finder_$f.json 2> gpurun_out/bench_finder_$f.err
  python -c "import json; d=json.load(open('gpurun_out/bench_finder_$f.json')); print('RUBIK finder $f', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1), d['config']['hit_finder'])" || tail -5 gpurun_out/bench_finder_$f.err
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --workload airplane-standin --finder $f > gpurun_out/bench_air_finder_$f.json 2> gpurun_out/bench_air_finder_$f.err
  python -c "import json; d=json.load(open('gpurun_out/bench_air_finder_$f.json')); print('AIR finder $f', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'], d['config']['hit_finder'])" || tail -5 gpurun_out/bench_air_finder_$f.err
done
SRT_WAVE_LOG=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu --finder 1 > /dev/null 2> gpurun_out/wave_log_finder.txt; head -14 gpurun_out/wave_log_finder.txt
