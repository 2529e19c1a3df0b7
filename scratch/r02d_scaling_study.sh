#!/bin/bash
# per-rank step of an 8-GPU band split emulated on one GPU (bench.py --stripe-of 8): where the time goes, and the tail threshold
cd "$(dirname "$0")/.."
for tp in -1 262144 131072 32768 0; do
  SRT_WAVE_LOG=1 python bench.py --stripe-of 8 --steps 3 --no-cpu --tail-paths $tp > gpurun_out/r02d_stripe8_tail$tp.json 2> gpurun_out/r02d_stripe8_tail$tp.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r02d_stripe8_tail$tp.json"))
print("stripe-of 8 tail_paths $tp: ms_per_step %.3f" % d["ms_per_step"])
PY
done
for so in 2 4; do
  SRT_WAVE_LOG=1 python bench.py --stripe-of $so --steps 3 --no-cpu > gpurun_out/r02d_stripe$so.json 2> gpurun_out/r02d_stripe$so.err
  python -c "import json;d=json.load(open('gpurun_out/r02d_stripe$so.json'));print('stripe-of $so: ms_per_step %.3f'%d['ms_per_step'])"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02d_stripe8_launches.csv python bench.py --stripe-of 8 --steps 2 --no-cpu > gpurun_out/r02d_ncu.log 2>&1
echo ncu rc $?
