#!/bin/bash
# round 3v: one fma per slab plane in the find kernel's descent (SRT_WF_FIND_FMA) -- parity, then A/B against the build without
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_adversarial_gpu.py tests/test_finder_build_gpu.py tests/test_fused_primary_gpu.py tests/test_checked_build_gpu.py -x -q -m gpu 2>&1 | tail -6
for rep in 1 2; do
for lib in "" scratch/variants/libsrt_nofma.so; do
  name=$([ -z "$lib" ] && echo fma || echo nofma)
  for wl in rubik airplane-standin; do
    SRT_LIB=$([ -z "$lib" ] || echo $PWD/$lib) timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu > gpurun_out/r03v_${name}_${wl}_$rep.json 2>gpurun_out/r03v_${name}_${wl}_$rep.err
    python -c "import json;d=json.load(open('gpurun_out/r03v_${name}_${wl}_$rep.json'));print('$name $wl $rep:', round(d['ms_per_step'],3),'ms', round(d['value'],1), 'trace ms', round(d['roofline']['kernel_ms_total'],2))"
  done
done
done
