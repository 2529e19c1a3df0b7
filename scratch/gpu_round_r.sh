#!/bin/bash
# round 1r: parity suite; launch list of the bench command; full ncu capture of the find / replay / logic kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01i_rubik.json 2> gpurun_out/bench_r01i_rubik.err
python -c "import json; d=json.load(open('gpurun_out/bench_r01i_rubik.json')); print('rubik', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['cpu_baseline'])" || tail -5 gpurun_out/bench_r01i_rubik.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_r.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01i_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1
tail -3 gpurun_out/r01i_launches.csv | cut -c1-200
timeout 300 python scratch/prof_one.py 3 1 32 0 > gpurun_out/plain_r2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"wf_find|wf_logic|wf_replay" -c 6 -o gpurun_out/prof_r01i python scratch/prof_one.py 3 1 32 0 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out | grep ncu-rep
