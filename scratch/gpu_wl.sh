#!/bin/bash
mkdir -p gpurun_out
for w in 8 4 1; do python scratch/wave_log.py 64 $w 2> gpurun_out/wl_$w.txt | tail -1; done
sed -n 97,108p gpurun_out/wl_8.txt
