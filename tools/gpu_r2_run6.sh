#!/bin/bash
# round 2, GPU run 6 (1 GPU; repeats what runs 4 and 5 measured -- their output was lost with the machine): engines agree at
# 8192x32768 after the lbm_create memset fix; L2 roof probe; fused seam barrier + same-GPU two-process ring; L2 eviction hints
# (speed and DRAM bytes); residency probe; two-chain probe; whole GPU suite
O=gpurun_out/r2_run6; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py --same-gpu --abort > $O/mgpu_same_gpu.log 2>&1; echo "same-gpu ring rc=$?" | tee -a $O/rc.txt; grep "\[mgpu\]" $O/mgpu_same_gpu.log | tail -3 | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_boundary.py tests/test_gpu_fullsize.py -m gpu -q -x > $O/pytest_quick.log 2>&1; echo "quick tests rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_quick.log
timeout 600 python tools/engine_diff.py dp 8192 32768 32 > $O/engine_diff_dp.log 2>&1; echo "diff dp rc=$?" | tee -a $O/rc.txt; tail -2 $O/engine_diff_dp.log | cut -c1-200
timeout 600 python tools/engine_diff.py sp 8192 32768 32 > $O/engine_diff_sp.log 2>&1; echo "diff sp rc=$?" | tee -a $O/rc.txt; tail -2 $O/engine_diff_sp.log | cut -c1-200
timeout 900 python tools/tile_sweep.py both hints > $O/hints_sweep.log 2>&1; echo "hints rc=$?" | tee -a $O/rc.txt; cat $O/hints_sweep.log | cut -c1-200
timeout 900 python tools/residency_probe.py both > $O/residency.log 2>&1; echo "residency rc=$?" | tee -a $O/rc.txt; cp gpurun_out/residency_probe.json $O/ 2>/dev/null
timeout 900 python tools/two_chain_probe.py dp > $O/two_chain_dp.log 2>&1; echo "two-chain dp rc=$?" | tee -a $O/rc.txt; cat $O/two_chain_dp.log | cut -c1-200
timeout 600 python tools/two_chain_probe.py sp > $O/two_chain_sp.log 2>&1; echo "two-chain sp rc=$?" | tee -a $O/rc.txt
timeout 900 python tools/l2_peak_probe.py > $O/l2_peak.log 2>&1; echo "l2 rc=$?" | tee -a $O/rc.txt; tail -1 $O/l2_peak.log; cp gpurun_out/l2_peak.json $O/ 2>/dev/null
for cfg in "dp large 32 inplace:3" "dp large 32 inplace:0" "sp large 32 inplace:3"; do
  tag=$(echo $cfg | tr ' :' '__')
  timeout 300 python tools/profile_target.py $cfg > $O/plain_$tag.log 2>&1 && \
  timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 20000 --csv --log-file $O/launches_$tag.csv python tools/profile_target.py $cfg > $O/ncu_l_$tag.log 2>&1
  cat $O/plain_$tag.log; wc -l $O/launches_$tag.csv
  set -- $cfg
  python tools/traffic_from_ncu.py $O/launches_$tag.csv ${1}_$(echo $4 | tr ':' 'h')_8192x32768 $1 8192 32768 32 > $O/traffic_$tag.txt 2>&1; tail -2 $O/traffic_$tag.txt | cut -c1-500
  gzip -f $O/launches_$tag.csv
done
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -8 $O/pytest_gpu.log
