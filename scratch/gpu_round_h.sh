#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -2 gpurun_out/pytest_gpu.log
show() { python -c "import json; d=json.load(open('gpurun_out/$1.json')); r=d['roofline']; print('$1', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'], 'frac', round(r['frac'],3), 'trace', round(r['kernel_ms_total']/d['steps'],2), 'logic', round(r['logic_kernel_ms_total']/d['steps'],2), 'n/t', round(r['nodes_per_ray'],1), round(r['tris_per_ray'],1), d['cpu_baseline'] and round(d['cpu_baseline']['value'],2))"; }
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/h_rubik.json 2> gpurun_out/h_rubik.err; show h_rubik
timeout 600 python bench.py --steps 3 --warmup 3 --workload airplane-standin --cpu-frames 1 > gpurun_out/h_airplane.json 2> gpurun_out/h_airplane.err; show h_airplane
timeout 900 python bench.py --steps 3 --warmup 3 --workload soup --no-cpu --l2-persist 1 > gpurun_out/h_soup_l2on.json 2> gpurun_out/h_soup_l2on.err; show h_soup_l2on
timeout 900 python bench.py --steps 3 --warmup 3 --workload soup --no-cpu --l2-persist 0 > gpurun_out/h_soup_l2off.json 2> gpurun_out/h_soup_l2off.err; show h_soup_l2off
tail -2 gpurun_out/h_soup_l2on.err
