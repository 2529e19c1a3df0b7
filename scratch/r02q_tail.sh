#!/bin/bash
# round 2q: tail threshold and per-rank step (emulated rank of 8) with the one-visit flow
cd "$(dirname "$0")/.."
for tp in 262144 524288 1048576 2097152; do
  for so in 1 8; do
    timeout 600 python bench.py --workload rubik --steps 10 --warmup 3 --no-cpu --tail-paths $tp --stripe-of $so > gpurun_out/r02q_tail${tp}_of$so.json 2> gpurun_out/r02q_tail${tp}_of$so.err
    python -c "import json;d=json.load(open('gpurun_out/r02q_tail${tp}_of$so.json'));print('tail $tp stripe-of $so:', round(d['ms_per_step'],3),'ms; launches/step', d['gpu_launches']/d['steps'])"
  done
done
for so in 2 4; do
  timeout 600 python bench.py --workload rubik --steps 10 --warmup 3 --no-cpu --stripe-of $so > gpurun_out/r02q_of$so.json 2> gpurun_out/r02q_of$so.err
  python -c "import json;d=json.load(open('gpurun_out/r02q_of$so.json'));print('default tail, stripe-of $so:', round(d['ms_per_step'],3),'ms')"
done
SRT_WAVE_LOG=1 timeout 600 python bench.py --workload rubik --steps 3 --warmup 3 --no-cpu --stripe-of 8 > /dev/null 2> gpurun_out/r02q_wavelog_of8.err; grep "^wave" gpurun_out/r02q_wavelog_of8.err | tail -9
