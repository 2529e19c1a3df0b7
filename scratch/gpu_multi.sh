#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?" >> gpurun_out/bench_n$N.err
tail -3 gpurun_out/bench_n$N.err
cat gpurun_out/bench_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N', d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'])"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
cat gpurun_out/bench_n1.json | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N', d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'])"
