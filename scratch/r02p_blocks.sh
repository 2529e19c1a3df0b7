#!/bin/bash
# round 2p: logic kernel block size (phases are block-wide): 128x8, 256x4 (default), 512x2, 1024x1
cd "$(dirname "$0")/.."
for v in default t128 t512 t1024; do
  if [ $v = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/scratch/variants/libsrt_$v.so; fi
  for w in rubik airplane-standin; do
    timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > gpurun_out/r02p_${w}_$v.json 2> gpurun_out/r02p_${w}_$v.err
    python -c "import json;d=json.load(open('gpurun_out/r02p_${w}_$v.json'));r=d['roofline'];print('$w $v:', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],3),'ms; logic ms/step', round(r.get('logic_kernel_ms_total',0)/d['steps'],2), 'ray kernels', round(r.get('kernel_ms_total',0)/d['steps'],2), 'e2e', round(d['e2e']['value'],1))"
  done
done
unset SRT_LIB
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02p_parity.log 2>&1; tail -3 gpurun_out/r02p_parity.log
