#!/bin/bash
# round 1m: tail kernel -- parity suite, then a sweep of the switch point on the bench workload
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for t in 0 262144 1048576 2097152 4194304 16777216; do
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --tail-paths $t > gpurun_out/bench_tail_$t.json 2> gpurun_out/bench_tail_$t.err
  python -c "import json; d=json.load(open('gpurun_out/bench_tail_$t.json')); print('TAIL $t', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1))"
done
SRT_WAVE_LOG=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu > /dev/null 2> gpurun_out/wave_log_tail.txt; head -14 gpurun_out/wave_log_tail.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --workload airplane-standin > gpurun_out/bench_air_tail.json 2> gpurun_out/bench_air_tail.err
python -c "import json; d=json.load(open('gpurun_out/bench_air_tail.json')); print('AIR', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'])"
