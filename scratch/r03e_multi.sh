#!/bin/bash
# round 3e (final build of round 2; r02l was the same script before the last kernel work): N GPUs of one box ($1 = N): Rubik 1080p/64 spp with both splits, the 10 M-triangle soup, (8 only) the stand-in at 4K/1024 spp
cd "$(dirname "$0")/.."
N=${1:-4}
run() {  # name, args...
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --no-cpu "$@" \
    > gpurun_out/r03e_${name}_${N}gpu.json 2> gpurun_out/r03e_${name}_${N}gpu.err
  python - gpurun_out/r03e_${name}_${N}gpu.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"], 1), "Mseg/s", round(d["ms_per_step"], 2), "ms; e2e", round(d["e2e"]["value"], 1), "; collective ms", d.get("collective_ms_per_step"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
run rubik_bands --workload rubik --steps 10 --warmup 3 --split bands
run rubik_samples --workload rubik --steps 10 --warmup 3 --split samples
run soup_bands --workload soup --steps 3 --warmup 3 --split bands
run standin_bands --workload airplane-standin --steps 5 --warmup 3 --split bands
if [ "$N" = "8" ]; then
  run standin4k_bands --workload airplane-standin-4k --steps 2 --warmup 3 --split bands
  run standin4k_samples --workload airplane-standin-4k --steps 2 --warmup 3 --split samples
fi
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r03e_pytest_multi_${N}gpu.log 2>&1; tail -3 gpurun_out/r03e_pytest_multi_${N}gpu.log
