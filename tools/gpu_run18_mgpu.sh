#!/bin/bash
# multi-GPU pass of the round-1 defaults (gpurun --gpus N): parity over the ring, bench.py exactly as the driver launches it
# (weak scaling, in-place engine + temporal blocking across the seams), strong scaling, and the A/B engine beside them
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L | head -8; free -g | head -2
echo "=== mgpu parity"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py > gpurun_out/mgpu_check_$N.log 2>&1
grep -c "BIT-EXACT" gpurun_out/mgpu_check_$N.log; grep -E "MISMATCH|Error" gpurun_out/mgpu_check_$N.log | head -5
echo "=== bench --gpus $N (driver's command: weak, in-place + temporal blocking)"
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err
cut -c1-250 gpurun_out/bench_$N.json; grep -v -i -E 'warn|OMP_NUM|\*\*\*|NCCL version|^$' gpurun_out/bench_$N.err | tail -5
echo "=== bench --gpus $N strong 16384x65536 DP in-place"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 3 --warmup 2 --workload strong --no-e2e > gpurun_out/bench_strong_$N.json 2> gpurun_out/bench_strong_$N.err
cut -c1-250 gpurun_out/bench_strong_$N.json; grep -v -i -E 'warn|OMP_NUM|\*\*\*|NCCL version|^$' gpurun_out/bench_strong_$N.err | tail -5
echo "=== bench --gpus $N weak A/B"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus $N --steps 2 --warmup 2 --engine ab --no-e2e > gpurun_out/bench_ab_$N.json 2> gpurun_out/bench_ab_$N.err
cut -c1-250 gpurun_out/bench_ab_$N.json; grep -v -i -E 'warn|OMP_NUM|\*\*\*|NCCL version|^$' gpurun_out/bench_ab_$N.err | tail -5
