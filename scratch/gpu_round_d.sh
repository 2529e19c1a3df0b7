#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for slots in 1048576 4194304 16777216; do
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --wave-slots $slots > gpurun_out/bench_wf_$slots.json 2> gpurun_out/bench_wf_$slots.err
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --block 128 > gpurun_out/bench_wf_b128.json 2> gpurun_out/bench_wf_b128.err
ls -la gpurun_out
