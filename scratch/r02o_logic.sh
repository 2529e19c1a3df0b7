#!/bin/bash
# round 2o: logic kernel in block phases, zero-numerator divisions, shared direct-light terms -- parity, then bench (64 / 80 registers)
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_adversarial_gpu.py tests/test_texture_option.py -m gpu -x -q > gpurun_out/r02o_parity.log 2>&1; tail -4 gpurun_out/r02o_parity.log
for v in default logic3; do
  if [ $v = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/scratch/variants/libsrt_$v.so; fi
  for w in rubik airplane-standin; do
    timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > gpurun_out/r02o_${w}_$v.json 2> gpurun_out/r02o_${w}_$v.err
    python -c "import json;d=json.load(open('gpurun_out/r02o_${w}_$v.json'));r=d['roofline'];print('$w $v:', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],3),'ms; logic ms/step', round(r.get('logic_kernel_ms_total',0)/d['steps'],2), 'ray kernels', round(r.get('kernel_ms_total',0)/d['steps'],2), 'e2e', round(d['e2e']['value'],1))"
  done
done
unset SRT_LIB
ncu --set full --import-source on --clock-control none -k regex:wf_logic_kernel -c 2 -o gpurun_out/r02o_logic python bench.py --workload rubik --steps 1 --warmup 0 --no-cpu > gpurun_out/r02o_ncu.log 2>&1; tail -2 gpurun_out/r02o_ncu.log
