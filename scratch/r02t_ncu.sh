#!/bin/bash
# round 2t: ncu metrics pass of one Rubik bench step (per-kernel table) + one --set full capture of every kernel of a step
cd "$(dirname "$0")/.."
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__cycles_elapsed.max
timeout 900 ncu --metrics $M --clock-control none -c 300 --csv --log-file gpurun_out/r02t_metrics_rubik.csv python bench.py --steps 2 --no-cpu > gpurun_out/r02t_ncu_metrics.log 2>&1; echo "ncu metrics rc $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wf_logic_kernel|wf_find_kernel|wf_replay_kernel|wf_tail_kernel|wf_init_kernel" -s 18 -c 12 -f -o gpurun_out/r02t_full python bench.py --steps 1 --no-cpu > gpurun_out/r02t_ncu_full.log 2>&1; echo "ncu full rc $?"; ls -la gpurun_out/r02t_full.ncu-rep
