#!/bin/bash
# round 2z: scene upload on the device end to end (re-layout + finder tables + finder trees + costing)
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_finder_build_gpu.py tests/test_device_relayout_gpu.py tests/test_parity_gpu.py tests/test_texture_option.py -m gpu -x -q > gpurun_out/r02z_pytest.log 2>&1; tail -4 gpurun_out/r02z_pytest.log
for w in airplane-standin rubik; do
  SRT_UPLOAD_LOG=1 SRT_BENCH_DEBUG=1 timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > gpurun_out/r02z_${w}.json 2> gpurun_out/r02z_${w}.err
  python -c "import json;d=json.load(open('gpurun_out/r02z_${w}.json'));e=d['e2e'];print('$w:', round(d['value'],1),'Mseg/s resident', round(d['ms_per_step'],2), 'ms;', round(e['value'],1), 'e2e', round(e['ms_per_step'],2),'ms;', e.get('scene_relayout'), d['config']['hit_finder'])"
  grep -E "srt_upload_scene|build_finder|e2e iteration" gpurun_out/r02z_${w}.err | tail -5
done
