#!/bin/bash
# round 1k: final state -- smoke, parity suite, default bench line (with the CPU baseline), reference arm, launch list, wave log, airplane
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r01k_bench_rubik.json 2> gpurun_out/r01k_bench_rubik.err
python -c "import json; d=json.load(open('gpurun_out/r01k_bench_rubik.json')); print('default', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['cpu_baseline']['value'], d['clocks'], d['roofline']['frac'])" || tail -5 gpurun_out/r01k_bench_rubik.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01k_bench_reference_arm.json 2> gpurun_out/r01k_ref.err; tail -c 300 gpurun_out/r01k_bench_reference_arm.json; echo
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --workload airplane-standin > gpurun_out/r01k_bench_airplane_standin.json 2> gpurun_out/r01k_air.err
python -c "import json; d=json.load(open('gpurun_out/r01k_bench_airplane_standin.json')); print('air', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1))"
SRT_WAVE_LOG=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu > /dev/null 2> gpurun_out/r01k_wave_log.txt; grep "^wave" gpurun_out/r01k_wave_log.txt | tail -10
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_k.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01k_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launches_k.log 2>&1; wc -l gpurun_out/r01k_launches.csv
