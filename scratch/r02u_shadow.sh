#!/bin/bash
# round 2u: shadow rays leave the finder walk at their first candidate
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_adversarial_gpu.py tests/test_finder_build_gpu.py tests/test_checked_build_gpu.py -m gpu -x -q > gpurun_out/r02u_parity.log 2>&1; tail -3 gpurun_out/r02u_parity.log
for w in rubik airplane-standin; do
  SRT_WAVE_LOG=1 timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > gpurun_out/r02u_$w.json 2> gpurun_out/r02u_$w.err
  python -c "import json;d=json.load(open('gpurun_out/r02u_$w.json'));r=d['roofline'];print('$w:', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],3),'ms; logic ms/step', round(r.get('logic_kernel_ms_total',0)/d['steps'],2), 'ray kernels', round(r.get('kernel_ms_total',0)/d['steps'],2), 'e2e', round(d['e2e']['value'],1), 'exec', r['executed_per_ray'])"
  grep "^wave" gpurun_out/r02u_$w.err | tail -7
done
timeout 600 python bench.py --workload rubik --steps 10 --warmup 3 --no-cpu --stripe-of 8 > gpurun_out/r02u_rubik_of8.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/r02u_rubik_of8.json'));print('rank of 8:', round(d['ms_per_step'],3),'ms')"
