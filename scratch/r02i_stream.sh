#!/bin/bash
# round 2i: slot records as streaming traffic in the logic/init kernels -- parity, step time, DRAM bytes per kernel
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > gpurun_out/r02i_pytest.log 2>&1; tail -2 gpurun_out/r02i_pytest.log
for w in rubik airplane-standin; do
  SRT_WAVE_LOG=1 python bench.py --workload $w --no-cpu > gpurun_out/r02i_bench_$w.json 2> gpurun_out/r02i_bench_$w.err
  python -c "import json;d=json.load(open('gpurun_out/r02i_bench_$w.json'));r=d['roofline'];print('$w', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],2),'ms; trace', round(r['kernel_ms_total']/d['steps'],2), 'logic', round(r['logic_kernel_ms_total']/d['steps'],2))"
done
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__cycles_elapsed.max
timeout 900 ncu --metrics $M --clock-control none -c 190 --csv --log-file gpurun_out/r02i_metrics_rubik.csv python bench.py --steps 2 --no-cpu > gpurun_out/r02i_ncu_metrics.log 2>&1; echo "ncu metrics rc $?"
