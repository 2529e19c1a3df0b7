#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python scratch/prof_one.py 3 1 16 > gpurun_out/plain_wf.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_wf.csv python scratch/prof_one.py 3 1 16 > gpurun_out/ncu_launches_wf.log 2>&1
ls -la gpurun_out
