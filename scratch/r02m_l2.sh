#!/bin/bash
# round 2m: L2 access-policy window on the noise tables (3) vs the BVH top levels (2) vs none (0); Rubik, stand-in, spheres
cd "$(dirname "$0")/.."
for w in rubik airplane-standin spheres; do
  for l2 in 0 2 3; do
    timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --l2-persist $l2 > gpurun_out/r02m_${w}_l2_$l2.json 2> gpurun_out/r02m_${w}_l2_$l2.err
    python -c "import json;d=json.load(open('gpurun_out/r02m_${w}_l2_$l2.json'));r=d['roofline'];print('$w l2=$l2:', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],3),'ms; logic ms/step', round(r.get('logic_kernel_ms_total',0)/d['steps'],2), 'ray kernels', round(r.get('kernel_ms_total',0)/d['steps'],2), 'e2e', round(d['e2e']['value'],1))"
  done
done
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_device_relayout_gpu.py -m gpu -x -q > gpurun_out/r02m_pytest.log 2>&1; tail -3 gpurun_out/r02m_pytest.log
