#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_wf.json 2> gpurun_out/bench_wf.err; echo "rc=$?" >> gpurun_out/bench_wf.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --schedule persistent > gpurun_out/bench_pers.json 2> gpurun_out/bench_pers.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --block 128 > gpurun_out/bench_wf128.json 2> gpurun_out/bench_wf128.err
ls -la gpurun_out
