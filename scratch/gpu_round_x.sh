#!/bin/bash
# round 1x: the 10 M-triangle soup (HBM-bound): bench line, wave log, full ncu capture of two trace launches
mkdir -p gpurun_out
SRT_WAVE_LOG=1 timeout 1200 python bench.py --workload soup --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_x_soup.json 2> gpurun_out/bench_x_soup.err
python -c "import json; d=json.load(open('gpurun_out/bench_x_soup.json')); print('soup', round(d['value'],2), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],2), d['gpu_launches'], d['roofline']['frac'], d['config']['hit_finder'])" || tail -5 gpurun_out/bench_x_soup.err
grep "^wave" gpurun_out/bench_x_soup.err | tail -40 | head -30
timeout 1800 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 20 -c 3 -o gpurun_out/prof_r01j_soup python bench.py --workload soup --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_x.log 2>&1; echo "ncu rc=$?"; ls -la gpurun_out | grep soup
