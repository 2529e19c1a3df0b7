#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/plain_k.json 2> gpurun_out/plain_k.err &&
timeout 1500 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum,l1tex__t_bytes.sum --clock-control none -k regex:wf_trace -c 400 --csv --log-file gpurun_out/trace_dram_r01f.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_k.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_k.log | cut -c1-300
