#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_bvh_build_gpu.py tests/test_parity_gpu.py -m gpu -x -q -k "bvh or two_models or tree or empty" > gpurun_out/pytest_bvh.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_bvh.log
tail -30 gpurun_out/pytest_bvh.log
