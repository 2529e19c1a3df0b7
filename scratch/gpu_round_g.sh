#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scratch/prof_one.py 3 1 16 > gpurun_out/plain_g.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 41 -c 2 -o gpurun_out/prof_r01c_trace python scratch/prof_one.py 3 1 16 > gpurun_out/ncu_g1.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:wf_logic -s 40 -c 2 -o gpurun_out/prof_r01c_logic python scratch/prof_one.py 3 1 16 > gpurun_out/ncu_g2.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 --workload airplane-standin --cpu-frames 1 > gpurun_out/bench_airplane.json 2> gpurun_out/bench_airplane.err
timeout 900 python bench.py --steps 3 --warmup 3 --workload soup --soup-tris 1000000 --cpu-frames 1 > gpurun_out/bench_soup1m.json 2> gpurun_out/bench_soup1m.err
ls -la gpurun_out | tail -8
