#!/bin/bash
# round 1w: smoke, parity suite, default bench line, reference arm, e2e per-iteration breakdown
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
SRT_BENCH_DEBUG=1 timeout 900 python bench.py > gpurun_out/bench_w_default.json 2> gpurun_out/bench_w_default.err
python -c "import json; d=json.load(open('gpurun_out/bench_w_default.json')); print('default', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['cpu_baseline'], d['clocks'])" || tail -5 gpurun_out/bench_w_default.err
grep "e2e iteration" gpurun_out/bench_w_default.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_w_reference.json 2> gpurun_out/bench_w_reference.err; tail -c 600 gpurun_out/bench_w_reference.json
