#!/bin/bash
# multi-GPU pass (gpurun --gpus N): parity over the ring, weak-scaling and strong-scaling bench lines
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8; free -g | head -2
echo "=== mgpu parity"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py > gpurun_out/mgpu_check_$N.log 2>&1
grep -E "mgpu|Error" gpurun_out/mgpu_check_$N.log | head -12
echo "=== C driver, $N GPUs in one process (8192 rows per GPU)"
timeout 300 ./driver/lbm_b200_dp $((8192*N)) 32768 200 $N 1 2>&1 | tail -3 | tee gpurun_out/driver_$N.log
echo "=== bench --gpus $N weak (p2p)"
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err
tail -c 2500 gpurun_out/bench_$N.json; grep -v -i -E 'warn|OMP_NUM|\*\*\*|^$' gpurun_out/bench_$N.err | tail -8
echo "=== bench --gpus $N strong 16384x65536 DP (p2p)"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 3 --warmup 2 --workload strong --no-e2e > gpurun_out/bench_strong_$N.json 2> gpurun_out/bench_strong_$N.err
tail -c 1800 gpurun_out/bench_strong_$N.json; grep -v -i -E 'warn|OMP_NUM|\*\*\*|^$' gpurun_out/bench_strong_$N.err | tail -8
if [ "$2" = "staged" ]; then
echo "=== bench --gpus $N weak (staged NCCL)"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 2 --warmup 1 --transport staged --no-e2e > gpurun_out/bench_staged_$N.json 2> gpurun_out/bench_staged_$N.err
tail -c 1200 gpurun_out/bench_staged_$N.json
fi
