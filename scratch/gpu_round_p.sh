#!/bin/bash
# round 1o: primary cull -- parity suite, cull on/off on Rubik + airplane stand-in, wave log, e2e breakdown
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
for c in 1; do
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --primary-cull $c > gpurun_out/bench_out_$c.json 2> gpurun_out/bench_out_$c.err
  python -c "import json; d=json.load(open('gpurun_out/bench_out_$c.json')); print('RUBIK cull $c', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1))" || tail -5 gpurun_out/bench_out_$c.err
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --workload airplane-standin --primary-cull $c > gpurun_out/bench_air_out_$c.json 2> gpurun_out/bench_air_out_$c.err
  python -c "import json; d=json.load(open('gpurun_out/bench_air_out_$c.json')); print('AIR cull $c', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'])" || tail -5 gpurun_out/bench_air_out_$c.err
done
SRT_WAVE_LOG=1 SRT_BENCH_DEBUG=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu > /dev/null 2> gpurun_out/wave_log_out.txt; head -16 gpurun_out/wave_log_out.txt
SRT_WAVE_LOG=1 SRT_BENCH_DEBUG=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu --workload airplane-standin > /dev/null 2> gpurun_out/wave_log_air_out.txt; head -30 gpurun_out/wave_log_air_out.txt
