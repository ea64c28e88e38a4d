#!/bin/bash
# multi-GPU pass (gpurun --gpus N): parity over the ring, then the weak-scaling bench line
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L; nvidia-smi topo -m | head -12; free -g | head -2
echo "=== mgpu parity"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py 2>&1 | grep -v -E 'Warning|warn' | tail -25 | tee gpurun_out/mgpu_check_$N.log
echo "=== C driver, $N GPUs in one process"
timeout 300 ./driver/lbm_b200_dp 8192 32768 200 $N 1 2>&1 | tail -4 | tee gpurun_out/driver_$N.log
echo "=== bench --gpus $N (p2p)"
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err
tail -c 3000 gpurun_out/bench_$N.json; grep -v -E 'Warning|warn' gpurun_out/bench_$N.err | tail -8
echo "=== bench --gpus $N (staged NCCL)"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 2 --warmup 1 --transport staged --no-e2e > gpurun_out/bench_staged_$N.json 2> gpurun_out/bench_staged_$N.err
tail -c 1500 gpurun_out/bench_staged_$N.json; grep -v -E 'Warning|warn' gpurun_out/bench_staged_$N.err | tail -8
