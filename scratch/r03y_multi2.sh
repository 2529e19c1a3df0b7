#!/bin/bash
# round 3y: 2 GPUs, the last build: multi-GPU tests and one Rubik line per split
cd "$(dirname "$0")/.."
N=2
timeout 200 python -m pytest tests/test_multi_gpu.py -q -m gpu 2>&1 | tail -3 | tee gpurun_out/r03y_pytest_multi_2gpu.log
run() {
  name=$1; shift
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --no-cpu "$@" \
    > gpurun_out/r03y_${name}_${N}gpu.json 2> gpurun_out/r03y_${name}_${N}gpu.err
  python - gpurun_out/r03y_${name}_${N}gpu.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["value"], 1), "Mseg/s", round(d["ms_per_step"], 2), "ms; e2e", round(d["e2e"]["value"], 1), "; collective ms", d.get("collective_ms_per_step"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
run rubik_samples --workload rubik --steps 10 --warmup 3
run rubik_bands --workload rubik --steps 10 --warmup 3 --split bands
