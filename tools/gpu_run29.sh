#!/bin/bash
# run 29: ncu of the fused pair kernel (why is it not bandwidth-bound?)
mkdir -p gpurun_out
python tools/profile_target.py dp small 12 pairs:8
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"pair_kernel|edges_kernel" -s 4 -c 2 -f -o gpurun_out/prof_dp_small_pairs python tools/profile_target.py dp small 12 pairs:8 > gpurun_out/ncu_pairs.log 2>&1; tail -2 gpurun_out/ncu_pairs.log
ncu -i gpurun_out/prof_dp_small_pairs.ncu-rep --page raw --csv > gpurun_out/prof_dp_small_pairs.raw.csv 2>/dev/null; wc -c gpurun_out/prof_dp_small_pairs.raw.csv
