#!/bin/bash
# run 17: round-1 verification on one GPU -- parity suite, smoke, bench (both arms), strong lattice in place, ncu of the temporally blocked large lattice
mkdir -p gpurun_out
echo "=== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu.log
echo "=== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tee gpurun_out/smoke.log
echo "=== bench"; timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print(d['value'], d['roofline']['frac'], d['e2e'], 'sp', d['sp']['value'], d['sp']['e2e']['value'], d['ab_double_buffer']['dp']['value'], d['ab_double_buffer']['sp']['value']); print(d['weak_scaling_base_1gpu']); print(d['clocks'], d['cpu_baseline'])"; tail -3 gpurun_out/bench.err
echo "=== ref arm"; timeout 600 python bench.py --impl reference --steps 2 --warmup 1 | tee gpurun_out/bench_ref.json | cut -c1-200
echo "=== strong 16384x65536 DP on one GPU (in place + 2-D temporal blocking)"
timeout 900 python bench.py --workload strong --steps 2 --warmup 1 > gpurun_out/bench_strong_1.json 2> gpurun_out/bench_strong_1.err; cut -c1-300 gpurun_out/bench_strong_1.json; tail -3 gpurun_out/bench_strong_1.err
echo "=== ncu"
for cfg in "dp large 16 inplace" "sp large 8 inplace"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 300 python tools/profile_target.py $cfg > gpurun_out/plain_$tag.log 2>&1 && \
  timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 4000 --csv --log-file gpurun_out/launches_$tag.csv python tools/profile_target.py $cfg > gpurun_out/ncu_l_$tag.log 2>&1
  cat gpurun_out/plain_$tag.log
done
timeout 300 python tools/profile_target.py dp large 16 inplace > /dev/null 2>&1 && \
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k regex:stream_collide_aa -s 1500 -c 2 -f -o gpurun_out/prof_dp_large_16_inplace_tb python tools/profile_target.py dp large 16 inplace > gpurun_out/ncu_f_dp_large_tb.log 2>&1; tail -1 gpurun_out/ncu_f_dp_large_tb.log
