#!/bin/bash
# round 1u: replay kernel occupancy variants, tail switch-point sweep, DRAM/L2/L1 bytes of the trace-step kernels
mkdir -p gpurun_out
run() {  # tag, extra args
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu $2 > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err
  python -c "import json; d=json.load(open('gpurun_out/bench_$1.json')); print('$1', round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), d['gpu_launches'], round(d['roofline']['kernel_ms_total'],1), round(d['roofline']['logic_kernel_ms_total'],1))" || tail -5 gpurun_out/bench_$1.err
}
run u_base ""
cp simple-ray-tracer_b200/libsrt_b200.so /tmp/keep.so
for mb in 6 8; do cp scratch/libsrt_replay$mb.so simple-ray-tracer_b200/libsrt_b200.so; run u_replay$mb ""; done
cp /tmp/keep.so simple-ray-tracer_b200/libsrt_b200.so
for t in 131072 262144 524288 2097152; do run u_tail_$t "--tail-paths $t"; done
run u_air "--workload airplane-standin"
SRT_UPLOAD_LOG=1 SRT_BENCH_DEBUG=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu --workload airplane-standin > /dev/null 2> gpurun_out/wave_log_air_u.txt; grep -v "^wave" gpurun_out/wave_log_air_u.txt | tail -3
timeout 1500 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum,l1tex__t_bytes.sum --clock-control none -k regex:"wf_find|wf_replay|wf_trace|wf_init|wf_tail" -c 600 --csv --log-file gpurun_out/r01i_trace_dram_bytes.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_u.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/r01i_trace_dram_bytes.csv
