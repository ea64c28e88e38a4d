#!/bin/bash
# round 2, multi-GPU pass (N = $1 GPUs): ring parity (17 cases + abort path at N = 2) and the driver's bench line with its strong sub-record
N=${1:-8}; O=gpurun_out/r2_run${RUN:-9}_${N}gpu; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tools/mgpu_check.py > $O/mgpu_check_$N.log 2>&1; echo "mgpu_check rc=$?" | tee -a $O/rc.txt; grep "\[mgpu\]" $O/mgpu_check_$N.log | tail -2 | cut -c1-250
timeout 1200 $TR --master-port 29513 bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_$N.json 2> $O/bench_$N.err; echo "bench rc=$?" | tee -a $O/rc.txt; tail -3 $O/bench_$N.err; cut -c1-600 $O/bench_$N.json
