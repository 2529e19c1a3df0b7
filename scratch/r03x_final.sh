#!/bin/bash
# round 3x: the last build -- whole GPU suite, bench line with CPU baseline, reference arm, stand-in line, ncu launch list (metrics pass)
cd "$(dirname "$0")/.."
timeout 400 python -m pytest tests -m gpu -q > gpurun_out/r03x_pytest_gpu.log 2>&1; tail -3 gpurun_out/r03x_pytest_gpu.log
python bench.py > gpurun_out/r03x_bench_rubik.json 2> gpurun_out/r03x_bench_rubik.err || tail -5 gpurun_out/r03x_bench_rubik.err
python -c "import json;d=json.load(open('gpurun_out/r03x_bench_rubik.json'));r=d['roofline'];print('rubik', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],2),'ms e2e',round(d['e2e']['value'],1), 'bound',r['bound'],'frac',round(r['frac'],3),'cpu',d['cpu_baseline'] and round(d['cpu_baseline']['value'],2))"
python bench.py --workload airplane-standin --no-cpu > gpurun_out/r03x_bench_standin.json 2>/dev/null
python -c "import json;d=json.load(open('gpurun_out/r03x_bench_standin.json'));r=d['roofline'];print('standin', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],2),'ms e2e',round(d['e2e']['value'],1), 'bound',r['bound'],'frac',round(r['frac'],3))"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__cycles_elapsed.max
timeout 300 ncu --metrics $M --clock-control none -c 200 --csv --log-file gpurun_out/r03x_metrics_rubik.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r03x_ncu_metrics.log 2>&1; echo "ncu metrics rc $?"
