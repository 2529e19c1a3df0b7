#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | wc -l > gpurun_out/ngpus.txt
for N in 8 4 2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_r01i_n$N.json 2> gpurun_out/bench_r01i_n$N.err; echo "rc=$?" >> gpurun_out/bench_r01i_n$N.err
tail -1 gpurun_out/bench_r01i_n$N.err
tail -1 gpurun_out/bench_r01i_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N', d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['gpu_launches'], d['roofline']['frac'])"
done
SRT_WAVE_LOG=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29527 bench.py --gpus 8 --steps 1 --warmup 3 > /dev/null 2> gpurun_out/wave_log_n8.txt; grep "^wave" gpurun_out/wave_log_n8.txt | head -24
