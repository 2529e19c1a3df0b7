#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_headless_gpu.py -m gpu -x -q > gpurun_out/pytest_headless.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_headless.log
tail -15 gpurun_out/pytest_headless.log
timeout 900 python scratch/bvh_bench.py 10000000 --cpu > gpurun_out/bvh_bench_10m.json 2> gpurun_out/bvh_bench.err; cat gpurun_out/bvh_bench_10m.json; tail -3 gpurun_out/bvh_bench.err
timeout 300 python scratch/bvh_bench.py 1000000 > gpurun_out/bvh_bench_1m.json 2>> gpurun_out/bvh_bench.err; cat gpurun_out/bvh_bench_1m.json
