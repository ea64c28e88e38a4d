#!/bin/bash
# round 2, GPU run 8 (2 GPUs): ring parity incl. the chain cases and the abort path over two real GPUs, slab sweep (strip width x
# chains x hints), the driver's multi-GPU bench line
O=gpurun_out/r2_run8_2gpu; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tools/mgpu_check.py --abort > $O/mgpu_check_2.log 2>&1; echo "mgpu_check rc=$?" | tee -a $O/rc.txt; grep "\[mgpu\]" $O/mgpu_check_2.log | tail -4 | cut -c1-250
timeout 600 $TR --master-port 29512 tools/mgpu_sweep.py both > $O/mgpu_sweep_2.log 2>&1; echo "sweep rc=$?" | tee -a $O/rc.txt; grep "^{" $O/mgpu_sweep_2.log | cut -c1-200
timeout 1200 $TR --master-port 29513 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_2.json 2> $O/bench_2.err; echo "bench rc=$?" | tee -a $O/rc.txt; tail -3 $O/bench_2.err; cut -c1-1500 $O/bench_2.json
