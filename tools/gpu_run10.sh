#!/bin/bash
# run 10: 2-D grid AA kernels + balanced bands; parity, finer tb sweep, bench with the in-place headline, ncu of the AA kernels
mkdir -p gpurun_out
echo "=== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu.log
echo "=== tb sweep"; timeout 900 python tools/tb_sweep.py small 2>&1 | tee gpurun_out/tb_sweep.log | grep -E "^best|took|rror"
echo "=== bench"; timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; cut -c1-1500 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
echo "=== ncu"
for cfg in "dp small 64 inplace" "sp small 12 inplace" "dp large 6 inplace"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 300 python tools/profile_target.py $cfg > gpurun_out/plain_$tag.log 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv python tools/profile_target.py $cfg > gpurun_out/ncu_l_$tag.log 2>&1
  timeout 300 python tools/profile_target.py $cfg > gpurun_out/plain2_$tag.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k regex:stream_collide_aa -s 40 -c 2 -f -o gpurun_out/prof_$tag python tools/profile_target.py $cfg > gpurun_out/ncu_f_$tag.log 2>&1
  cat gpurun_out/plain_$tag.log; tail -1 gpurun_out/ncu_f_$tag.log
done
