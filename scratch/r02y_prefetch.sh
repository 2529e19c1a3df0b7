#!/bin/bash
# round 2y: L2 prefetch of the records phase B of the logic kernel gathers (on / off)
cd "$(dirname "$0")/.."
for v in default nopf default nopf; do
  if [ $v = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/scratch/variants/libsrt_$v.so; fi
  for w in rubik airplane-standin; do
    timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > gpurun_out/r02y_${w}_$v.json 2>/dev/null
    python -c "import json;d=json.load(open('gpurun_out/r02y_${w}_$v.json'));r=d['roofline'];print('$w $v:', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],3),'ms; logic ms/step', round(r.get('logic_kernel_ms_total',0)/d['steps'],2))"
  done
done
unset SRT_LIB
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -2
