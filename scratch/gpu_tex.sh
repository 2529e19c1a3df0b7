#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_texture_option.py tests/test_headless_gpu.py -x -q > gpurun_out/pytest_tex.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tex.log
tail -30 gpurun_out/pytest_tex.log
