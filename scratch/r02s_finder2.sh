#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_finder_build_gpu.py -m gpu -x -q > gpurun_out/r02s_pytest.log 2>&1; tail -3 gpurun_out/r02s_pytest.log
for w in airplane-standin rubik; do
  SRT_UPLOAD_LOG=1 timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/r02s_${w}.json 2> gpurun_out/r02s_${w}.err
  python -c "import json;d=json.load(open('gpurun_out/r02s_${w}.json'));e=d['e2e'];print('$w:', round(d['value'],1),'Mseg/s resident', round(d['ms_per_step'],2), 'ms;', round(e['value'],1), 'e2e', round(e['ms_per_step'],2),'ms')"
  grep -E "srt_upload_scene|build_finder" gpurun_out/r02s_${w}.err | tail -4
done
