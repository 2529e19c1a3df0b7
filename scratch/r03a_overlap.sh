#!/bin/bash
# round 3a (still round 2 of the build): SRT_OPT_OVERLAP -- tests, then bench on / off, 1 GPU and an emulated rank of 8
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_overlap_gpu.py -m gpu -x -q > gpurun_out/r03a_overlap_pytest.log 2>&1; tail -5 gpurun_out/r03a_overlap_pytest.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r03a_pytest.log 2>&1; tail -3 gpurun_out/r03a_pytest.log
for ov in 1 0; do
  for so in 1 8; do
    timeout 600 python bench.py --workload rubik --steps 10 --warmup 3 --no-cpu --overlap $ov --stripe-of $so > gpurun_out/r03a_rubik_ov${ov}_of$so.json 2> gpurun_out/r03a_rubik_ov${ov}_of$so.err
    python -c "import json;d=json.load(open('gpurun_out/r03a_rubik_ov${ov}_of$so.json'));print('rubik overlap $ov stripe-of $so:', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],3),'ms; e2e', round(d['e2e']['value'],1))"
  done
done
timeout 600 python bench.py --workload airplane-standin --steps 10 --warmup 3 --no-cpu --overlap 1 > gpurun_out/r03a_standin_ov1.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/r03a_standin_ov1.json'));print('standin overlap 1:', round(d['value'],1), round(d['ms_per_step'],3))"
