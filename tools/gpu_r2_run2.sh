#!/bin/bash
# round 2, GPU run 2: full-size parity tests, L2 roof probe (own kernels), restructured bench, ncu launch lists for traffic.json
O=gpurun_out/r2_run2; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_fullsize.py -x -q > $O/fullsize_test.log 2>&1; echo "fullsize rc=$?" | tee -a $O/rc.txt; tail -15 $O/fullsize_test.log
timeout 600 python tools/l2_peak_probe.py > $O/l2_peak.log 2>&1; echo "l2 rc=$?" | tee -a $O/rc.txt; tail -8 $O/l2_peak.log; cp gpurun_out/l2_peak.json $O/ 2>/dev/null
timeout 1500 python bench.py --steps 3 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt; tail -3 $O/bench.err; cut -c1-1200 $O/bench.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt; cut -c1-600 $O/bench_ref.json
for cfg in "dp large 32 inplace" "sp large 32 inplace" "dp small 64 inplace" "dp c3 64 inplace"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 300 python tools/profile_target.py $cfg > $O/plain_$tag.log 2>&1 && \
  timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 20000 --csv --log-file $O/launches_$tag.csv python tools/profile_target.py $cfg > $O/ncu_l_$tag.log 2>&1
  cat $O/plain_$tag.log; wc -l $O/launches_$tag.csv
done
