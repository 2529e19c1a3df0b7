#!/bin/bash
# round 2r: finder trees built on the device -- tests, then the stand-in and Rubik end to end with the build on the device / on host threads
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_finder_build_gpu.py -m gpu -x -q > gpurun_out/r02r_pytest.log 2>&1; tail -4 gpurun_out/r02r_pytest.log
for w in airplane-standin rubik; do
  for fb in 1 0; do
    SRT_UPLOAD_LOG=1 SRT_BENCH_DEBUG=1 timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --finder-build $fb > gpurun_out/r02r_${w}_fb$fb.json 2> gpurun_out/r02r_${w}_fb$fb.err
    python -c "import json;d=json.load(open('gpurun_out/r02r_${w}_fb$fb.json'));e=d['e2e'];print('$w finder build on', 'device' if $fb else 'host', ':', round(d['value'],1),'Mseg/s resident', round(d['ms_per_step'],2), 'ms;', round(e['value'],1), 'e2e', round(e['ms_per_step'],2),'ms; finder', d['config']['hit_finder'])"
    grep -E "srt_upload_scene|build_tree|build_finder" gpurun_out/r02r_${w}_fb$fb.err | tail -3
  done
done
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02r_pytest_all.log 2>&1; tail -4 gpurun_out/r02r_pytest_all.log
