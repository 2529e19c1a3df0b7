#!/bin/bash
# round 2n: one logic visit per path and wave (two rays per slot in flight) -- parity first (checked build too), then the bench lines
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_adversarial_gpu.py -m gpu -x -q > gpurun_out/r02n_parity.log 2>&1; tail -4 gpurun_out/r02n_parity.log
timeout 600 python -m pytest tests/test_checked_build_gpu.py -m gpu -x -q > gpurun_out/r02n_checked.log 2>&1; tail -3 gpurun_out/r02n_checked.log
for v in default logic3 logic2; do
  if [ $v = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/scratch/variants/libsrt_$v.so; fi
  timeout 600 python bench.py --workload rubik --steps 10 --warmup 3 --no-cpu > gpurun_out/r02n_rubik_$v.json 2> gpurun_out/r02n_rubik_$v.err
  python -c "import json;d=json.load(open('gpurun_out/r02n_rubik_$v.json'));r=d['roofline'];print('rubik $v:', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],3),'ms; logic ms/step', round(r.get('logic_kernel_ms_total',0)/d['steps'],2), 'ray kernels', round(r.get('kernel_ms_total',0)/d['steps'],2), 'e2e', round(d['e2e']['value'],1), 'launches/step', d['gpu_launches']/d['steps'])"
done
unset SRT_LIB
ncu --set full --import-source on --clock-control none -k regex:wf_logic_kernel -c 4 -o gpurun_out/r02n_logic python bench.py --workload rubik --steps 1 --warmup 0 --no-cpu > gpurun_out/r02n_ncu.log 2>&1; tail -2 gpurun_out/r02n_ncu.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02n_pytest.log 2>&1; tail -4 gpurun_out/r02n_pytest.log
