#!/bin/bash
# round 1y: culled-path bit map, L2 prefetch on the soup, ncu capture of the tail kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
run() {  # tag, extra args
  timeout 1200 python bench.py --steps $3 --warmup 3 --no-cpu $2 > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err
  python -c "import json; d=json.load(open('gpurun_out/bench_$1.json')); print('$1', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1), round(d['roofline']['frac'],3))" || tail -5 gpurun_out/bench_$1.err
}
run y_rubik "" 5
run y_air "--workload airplane-standin" 3
run y_soup_pf0 "--workload soup --prefetch 0" 2
run y_soup_pf1 "--workload soup --prefetch 1" 2
run y_rubik_pf1 "--finder 0 --prefetch 1" 3
run y_rubik_pf0 "--finder 0 --prefetch 0" 3
timeout 300 python scratch/prof_one.py 3 1 32 0 > gpurun_out/plain_y.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"wf_tail|wf_init|wf_resolve" -c 3 -o gpurun_out/prof_r01j_tail python scratch/prof_one.py 3 1 32 0 > gpurun_out/ncu_y.log 2>&1; echo "ncu rc=$?"
