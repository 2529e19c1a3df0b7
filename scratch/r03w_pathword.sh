#!/bin/bash
# round 3w: the path's number in the free word of the throughput record (no 8-byte record of its own) -- parity, then A/B
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_overlap_gpu.py tests/test_fused_primary_gpu.py tests/test_checked_build_gpu.py -x -q -m gpu 2>&1 | tail -6
for rep in 1 2; do
for lib in "" scratch/variants/libsrt_base.so; do
  name=$([ -z "$lib" ] && echo new || echo base)
  for wl in rubik airplane-standin; do
    SRT_LIB=$([ -z "$lib" ] || echo $PWD/$lib) timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu > gpurun_out/r03w_${name}_${wl}_$rep.json 2>gpurun_out/r03w_${name}_${wl}_$rep.err
    python -c "import json;d=json.load(open('gpurun_out/r03w_${name}_${wl}_$rep.json'));print('$name $wl $rep:', round(d['ms_per_step'],3),'ms', round(d['value'],1), 'logic ms', round(d['roofline']['logic_kernel_ms_total'],2))"
  done
done
done
