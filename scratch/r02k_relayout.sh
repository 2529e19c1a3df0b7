#!/bin/bash
# round 2k: device re-layout -- its tests, then the soup end to end with the re-layout on the device / host
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_device_relayout_gpu.py tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02k_relayout_pytest.log 2>&1; tail -5 gpurun_out/r02k_relayout_pytest.log
for dr in 2; do
  SRT_UPLOAD_LOG=1 SRT_BENCH_DEBUG=1 timeout 900 python bench.py --workload soup --steps 4 --no-cpu --device-relayout $dr > gpurun_out/r02k_soup_dr$dr.json 2> gpurun_out/r02k_soup_dr$dr.err
  python -c "import json;d=json.load(open('gpurun_out/r02k_soup_dr$dr.json'));e=d['e2e'];print('soup relayout $dr:', round(d['value'],1),'Mseg/s resident;', round(e['value'],1), 'e2e,', round(e['ms_per_step'],1),'ms/step;', e.get('scene_relayout'))"
  grep -E "e2e iteration|srt_upload_scene" gpurun_out/r02k_soup_dr$dr.err | tail -12
done
