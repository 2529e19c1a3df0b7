#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scratch/prof_one.py 3 1 32 0 > gpurun_out/plain_j.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 1 -c 2 -o gpurun_out/prof_r01f_trace python scratch/prof_one.py 3 1 32 0 > gpurun_out/ncu_j1.log 2>&1
ls -la gpurun_out | tail -4
