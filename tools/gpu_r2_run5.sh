#!/bin/bash
# round 2, GPU run 5 (1 GPU): fused seam barrier + same-GPU two-process ring, L2 eviction-priority hints (speed and DRAM bytes),
# residency probe (smem_pad), whole GPU suite with the new build
O=gpurun_out/r2_run5; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_boundary.py tests/test_gpu_fullsize.py -m gpu -q -x > $O/pytest_seam.log 2>&1; echo "seam tests rc=$?" | tee -a $O/rc.txt; tail -5 $O/pytest_seam.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py --same-gpu --abort > $O/mgpu_same_gpu.log 2>&1; echo "same-gpu ring rc=$?" | tee -a $O/rc.txt; tail -4 $O/mgpu_same_gpu.log | cut -c1-300
timeout 900 python tools/tile_sweep.py both hints > $O/hints_sweep.log 2>&1; echo "hints rc=$?" | tee -a $O/rc.txt; cat $O/hints_sweep.log | cut -c1-200
timeout 900 python tools/residency_probe.py both > $O/residency.log 2>&1; echo "residency rc=$?" | tee -a $O/rc.txt; cp gpurun_out/residency_probe.json $O/ 2>/dev/null; sort -t: -k9 $O/residency.log | tail -5
for cfg in "dp large 32 inplace:3" "sp large 32 inplace:3" "dp large 32 inplace:2"; do
  tag=$(echo $cfg | tr ' :' '__')
  timeout 300 python tools/profile_target.py $cfg > $O/plain_$tag.log 2>&1 && \
  timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -c 20000 --csv --log-file $O/launches_$tag.csv python tools/profile_target.py $cfg > $O/ncu_l_$tag.log 2>&1
  cat $O/plain_$tag.log; wc -l $O/launches_$tag.csv
  set -- $cfg
  python tools/traffic_from_ncu.py $O/launches_$tag.csv ${1}_$(echo $4 | tr ':' '_h')_8192x32768 $1 8192 32768 32 > $O/traffic_$tag.txt 2>&1; tail -4 $O/traffic_$tag.txt | cut -c1-400
  gzip -f $O/launches_$tag.csv
done
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -8 $O/pytest_gpu.log
