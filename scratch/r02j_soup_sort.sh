#!/bin/bash
# round 2j: ray queues sorted by origin cell -- parity, then the soup with the sort off / on
cd "$(dirname "$0")/.."
python -m pytest tests/test_parity_gpu.py tests/test_checked_build_gpu.py -m gpu -x -q -k "sorted or checked" > gpurun_out/r02j_pytest.log 2>&1; tail -3 gpurun_out/r02j_pytest.log
for sr in 0 1; do
  SRT_WAVE_LOG=1 python bench.py --workload soup --steps 3 --no-cpu --sort-rays $sr > gpurun_out/r02j_soup_sort$sr.json 2> gpurun_out/r02j_soup_sort$sr.err
  python -c "import json;d=json.load(open('gpurun_out/r02j_soup_sort$sr.json'));r=d['roofline'];print('soup sort $sr:', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],2),'ms; frac',round(r['frac'],3))"
  grep '^wave' gpurun_out/r02j_soup_sort$sr.err | tail -8 | awk '{printf "%s ", $4} END {print ""}'
done
