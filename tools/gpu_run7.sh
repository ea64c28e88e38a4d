#!/bin/bash
mkdir -p gpurun_out
echo "=== l2 probe"; timeout 600 python tools/l2_probe.py 2>&1 | tee gpurun_out/l2_probe.log
echo "=== bench"; timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 1200 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
echo "=== ref arm"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 | tee gpurun_out/bench_ref.json | cut -c1-300
echo "=== ncu"
for cfg in "dp small 12" "sp small 12" "dp large 6" "sp large 6"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 300 python tools/profile_target.py $cfg > gpurun_out/plain_$tag.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$tag.csv python tools/profile_target.py $cfg > gpurun_out/ncu_l_$tag.log 2>&1
  timeout 300 python tools/profile_target.py $cfg > gpurun_out/plain2_$tag.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_collide_vec -s 4 -c 2 -f -o gpurun_out/prof_$tag python tools/profile_target.py $cfg > gpurun_out/ncu_f_$tag.log 2>&1
  cat gpurun_out/plain_$tag.log; tail -1 gpurun_out/ncu_f_$tag.log
done
