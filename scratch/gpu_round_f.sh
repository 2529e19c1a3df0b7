#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scratch/prof_one.py 3 1 16 > gpurun_out/plain_f.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 3 -c 2 -o gpurun_out/prof_r01b_trace python scratch/prof_one.py 3 1 16 > gpurun_out/ncu_f1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_logic -s 3 -c 2 -o gpurun_out/prof_r01b_logic python scratch/prof_one.py 3 1 16 > gpurun_out/ncu_f2.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01b.csv python scratch/prof_one.py 3 1 16 > gpurun_out/ncu_f3.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01b.json 2> gpurun_out/bench_r01b.err
ls -la gpurun_out | tail -8
