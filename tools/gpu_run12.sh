#!/bin/bash
# run 12: register-cap experiment for the DP in-place kernels, strong-scaling lattice on one GPU (in place), ncu of SP small / DP large in place
mkdir -p gpurun_out
echo "=== occupancy probe"
for d in "" mb10 mb12 mb16; do
  if [ -z "$d" ]; then timeout 300 python tools/occ_probe.py; else LBM_B200_LIBDIR=$PWD/latticeboltzmann_b200/exp/$d timeout 300 python tools/occ_probe.py; fi
done 2>&1 | tee gpurun_out/occ_probe.log
echo "=== strong 16384x65536 DP on one GPU (in place)"
timeout 900 python bench.py --workload strong --steps 2 --warmup 1 > gpurun_out/bench_strong_1.json 2> gpurun_out/bench_strong_1.err; cut -c1-400 gpurun_out/bench_strong_1.json; tail -3 gpurun_out/bench_strong_1.err
echo "=== ncu"
for cfg in "sp small 12 inplace" "dp large 6 inplace"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 300 python tools/profile_target.py $cfg > gpurun_out/plain2_$tag.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_collide_aa -s 4 -c 2 -f -o gpurun_out/prof_$tag python tools/profile_target.py $cfg > gpurun_out/ncu_f_$tag.log 2>&1
  cat gpurun_out/plain2_$tag.log; tail -1 gpurun_out/ncu_f_$tag.log
done
