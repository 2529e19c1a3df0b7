#!/bin/bash
# round 2v: tail threshold on the stand-in (its reference BVH has leaves of up to 165 triangles: the tail's walk is slow there)
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02v_parity.log 2>&1; tail -2 gpurun_out/r02v_parity.log
for tp in 0 32768 131072 262144 524288 1048576; do
  for w in airplane-standin rubik; do
    timeout 600 python bench.py --workload $w --steps 6 --warmup 3 --no-cpu --tail-paths $tp > gpurun_out/r02v_${w}_$tp.json 2>/dev/null
    python -c "import json;d=json.load(open('gpurun_out/r02v_${w}_$tp.json'));print('$w tail $tp:', round(d['ms_per_step'],3),'ms; launches/step', d['gpu_launches']/d['steps'])"
  done
done
