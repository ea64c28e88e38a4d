#!/bin/bash
# round 2, GPU run 11 (1 GPU): verification of the final build -- smoke, whole GPU suite, bench (both arms), ncu launch lists of the
# default engine for profiles/traffic.json, one ncu --set full capture of its two kernels, hints in a long power-capped run
O=gpurun_out/r2_run11; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/gpu.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt; tail -2 $O/smoke.log | cut -c1-200
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_gpu.log
timeout 1500 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt; tail -2 $O/bench.err | cut -c1-200; cut -c1-400 $O/bench.json
timeout 900 python bench.py --impl reference > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt; cut -c1-500 $O/bench_ref.json
timeout 600 python tools/long_hints_probe.py both > $O/long_hints.log 2>&1; echo "long hints rc=$?" | tee -a $O/rc.txt; cat $O/long_hints.log
for cfg in "dp large 32 inplace" "sp large 32 inplace" "dp small 64 inplace" "dp c3 64 inplace"; do
  tag=$(echo $cfg | tr ' :' '__')
  timeout 300 python tools/profile_target.py $cfg > $O/plain_$tag.log 2>&1 && \
  timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 30000 --csv --log-file $O/launches_$tag.csv python tools/profile_target.py $cfg > $O/ncu_l_$tag.log 2>&1
  cat $O/plain_$tag.log; wc -l $O/launches_$tag.csv
  gzip -f $O/launches_$tag.csv
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_collide_aa -s 3000 -c 6 -f -o $O/prof_dp_large_inplace python tools/profile_target.py dp large 16 inplace > $O/ncu_full_dp.log 2>&1; echo "ncu full dp rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_collide_aa -s 1500 -c 6 -f -o $O/prof_sp_large_inplace python tools/profile_target.py sp large 16 inplace > $O/ncu_full_sp.log 2>&1; echo "ncu full sp rc=$?" | tee -a $O/rc.txt
ls -la $O
