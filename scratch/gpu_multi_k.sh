#!/bin/bash
mkdir -p gpurun_out
for N in 8 4 2; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_r01k_n$N.json 2> gpurun_out/bench_r01k_n$N.err; echo "rc=$?" >> gpurun_out/bench_r01k_n$N.err
tail -1 gpurun_out/bench_r01k_n$N.err
tail -1 gpurun_out/bench_r01k_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N', d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['gpu_launches'], d['roofline']['frac'])"
done
