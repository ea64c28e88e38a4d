#!/bin/bash
# round 2, GPU run 12 (1 GPU): last check of the shipped defaults (DP large lattices now with hints = 3): whole GPU suite, smoke, bench,
# ncu launch list of the DP headline for profiles/traffic.json
O=gpurun_out/r2_run12; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 1500 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt; cut -c1-300 $O/bench.json
timeout 600 python tools/engine_diff.py dp 8192 32768 32 > $O/engine_diff_dp.log 2>&1; echo "diff dp rc=$?" | tee -a $O/rc.txt; tail -1 $O/engine_diff_dp.log
cfg="dp large 32 inplace"; tag=dp_large_32_inplace
timeout 300 python tools/profile_target.py $cfg > $O/plain_$tag.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 30000 --csv --log-file $O/launches_$tag.csv python tools/profile_target.py $cfg > $O/ncu_l_$tag.log 2>&1
cat $O/plain_$tag.log; gzip -f $O/launches_$tag.csv
