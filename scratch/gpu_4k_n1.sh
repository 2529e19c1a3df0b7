#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 2 --warmup 3 --workload airplane-standin-4k --no-cpu > gpurun_out/bench_4k_n1.json 2> gpurun_out/bench_4k_n1.err; echo "rc=$?" >> gpurun_out/bench_4k_n1.err
tail -2 gpurun_out/bench_4k_n1.err
tail -1 gpurun_out/bench_4k_n1.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N', d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'], d['roofline']['frac'])"
