#!/bin/bash
# multi-GPU pass for the in-place engine (gpurun --gpus N): ring parity incl. in-place slabs, weak and strong bench lines
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
echo "=== mgpu parity (A/B and in-place slabs)"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py > gpurun_out/mgpu_check_inplace_$N.log 2>&1
grep -E "mgpu|Error" gpurun_out/mgpu_check_inplace_$N.log | head -14
echo "=== C driver in place, $N GPUs in one process (8192 rows per GPU)"
timeout 300 ./driver/lbm_b200_dp $((8192*N)) 32768 200 $N 1 0 1 2>&1 | tail -3 | tee gpurun_out/driver_inplace_$N.log
echo "=== bench --gpus $N weak in-place (p2p)"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --engine inplace --no-e2e > gpurun_out/bench_inplace_$N.json 2> gpurun_out/bench_inplace_$N.err
cut -c1-300 gpurun_out/bench_inplace_$N.json; grep -v -i -E 'warn|OMP_NUM|\*\*\*|^$' gpurun_out/bench_inplace_$N.err | tail -5
echo "=== bench --gpus $N strong 16384x65536 DP in-place"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 3 --warmup 2 --workload strong --engine inplace --no-e2e > gpurun_out/bench_strong_inplace_$N.json 2> gpurun_out/bench_strong_inplace_$N.err
cut -c1-300 gpurun_out/bench_strong_inplace_$N.json; grep -v -i -E 'warn|OMP_NUM|\*\*\*|^$' gpurun_out/bench_strong_inplace_$N.err | tail -5
echo "=== bench --gpus $N strong 16384x65536 DP A/B"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus $N --steps 3 --warmup 2 --workload strong --no-e2e > gpurun_out/bench_strong_$N.json 2> gpurun_out/bench_strong_$N.err
cut -c1-300 gpurun_out/bench_strong_$N.json; grep -v -i -E 'warn|OMP_NUM|\*\*\*|^$' gpurun_out/bench_strong_$N.err | tail -5
