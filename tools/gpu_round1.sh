#!/bin/bash
# first GPU pass: tests, bench, sweep, ncu (one box, one GPU)
mkdir -p gpurun_out
nvidia-smi -L; free -g | head -2; nproc; lscpu | grep -E 'Model name|Socket|Core|Thread' 
echo "=== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
echo "=== pytest gpu"; timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -40 | tee gpurun_out/pytest_gpu.log
echo "=== bench"; timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
echo "=== sweep"; timeout 900 python tools/sweep.py > gpurun_out/sweep.log 2>&1; tail -45 gpurun_out/sweep.log
echo "=== ncu"
for cfg in "dp small 20" "dp large 6" "sp large 6"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 300 python tools/profile_target.py $cfg > gpurun_out/plain_$tag.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$tag.csv python tools/profile_target.py $cfg > gpurun_out/ncu_l_$tag.log 2>&1
  timeout 300 python tools/profile_target.py $cfg > gpurun_out/plain2_$tag.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_collide_vec -s 4 -c 2 -f -o gpurun_out/prof_$tag python tools/profile_target.py $cfg > gpurun_out/ncu_f_$tag.log 2>&1
  cat gpurun_out/plain_$tag.log; tail -2 gpurun_out/ncu_f_$tag.log
done
ls -la gpurun_out
