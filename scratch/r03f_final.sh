#!/bin/bash
# round 3f: final build of round 2 (after SRT_OPT_OVERLAP) -- whole GPU suite, checked-build driver, bench lines of every workload, ncu metrics pass + one --set full capture (Rubik)
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r03f_pytest_gpu.log 2>&1; tail -3 gpurun_out/r03f_pytest_gpu.log
python tests/kernel_driver.py > gpurun_out/r03f_driver_normal.log 2>&1
SRT_LIB=$PWD/simple-ray-tracer_b200/libsrt_b200_checked.so python tests/kernel_driver.py > gpurun_out/r03f_driver_checked.log 2>&1; tail -2 gpurun_out/r03f_driver_checked.log
for w in rubik spheres airplane-standin; do
  python bench.py --workload $w > gpurun_out/r03f_bench_$w.json 2> gpurun_out/r03f_bench_$w.err || tail -5 gpurun_out/r03f_bench_$w.err
  python -c "import json;d=json.load(open('gpurun_out/r03f_bench_$w.json'));r=d['roofline'];print('$w', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],2),'ms e2e',round(d['e2e']['value'],1), 'bound',r['bound'],'frac',round(r['frac'],3),'issue',round(r.get('frac_issue',0),3),'cpu',d['cpu_baseline'] and round(d['cpu_baseline']['value'],2))"
done
python bench.py --workload soup --steps 3 > gpurun_out/r03f_bench_soup.json 2> gpurun_out/r03f_bench_soup.err || tail -5 gpurun_out/r03f_bench_soup.err
python -c "import json;d=json.load(open('gpurun_out/r03f_bench_soup.json'));r=d['roofline'];print('soup', round(d['value'],1),'Mseg/s', round(d['ms_per_step'],2),'ms e2e',round(d['e2e']['value'],1), 'bound',r['bound'],'frac',round(r['frac'],3))"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r03f_bench_reference.json 2>/dev/null; cut -c1-400 gpurun_out/r03f_bench_reference.json
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__cycles_elapsed.max
timeout 900 ncu --metrics $M --clock-control none -c 300 --csv --log-file gpurun_out/r03f_metrics_rubik.csv python bench.py --steps 2 --no-cpu > gpurun_out/r03f_ncu_metrics.log 2>&1; echo "ncu metrics rc $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wf_logic_kernel|wf_find_kernel|wf_replay_kernel|wf_tail_kernel|wf_init_kernel" -s 18 -c 17 -f -o gpurun_out/r03f_full python bench.py --steps 1 --no-cpu > gpurun_out/r03f_ncu_full.log 2>&1; echo "ncu full rc $?"; ls -la gpurun_out/r03f_full.ncu-rep
for tp in 524288 2097152 4194304; do
  for so in 1 8; do
    timeout 600 python bench.py --workload rubik --steps 10 --warmup 3 --no-cpu --tail-paths $tp --stripe-of $so > gpurun_out/r03f_tail${tp}_of$so.json 2>/dev/null
    python -c "import json;d=json.load(open('gpurun_out/r03f_tail${tp}_of$so.json'));print('overlap on, tail $tp stripe-of $so:', round(d['ms_per_step'],3),'ms')"
  done
done
for w in rubik airplane-standin; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --stripe-of 8 > gpurun_out/r03f_${w}_of8.json 2>/dev/null
  python -c "import json;d=json.load(open('gpurun_out/r03f_${w}_of8.json'));print('$w, emulated rank of 8:', round(d['ms_per_step'],3),'ms')"
done
