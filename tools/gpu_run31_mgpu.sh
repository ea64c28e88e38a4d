#!/bin/bash
# run 31 (gpurun --gpus 2): ring parity and weak-scaling bench of the final round-1 build (SP kernels at 64 registers, SP depth 16)
N=${1:-2}
mkdir -p gpurun_out
echo "=== mgpu parity"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py > gpurun_out/mgpu_check_$N.log 2>&1
grep -c "BIT-EXACT" gpurun_out/mgpu_check_$N.log; grep -E "MISMATCH|Error" gpurun_out/mgpu_check_$N.log | head -5
echo "=== bench --gpus $N (weak, in-place + temporal blocking, no e2e leg)"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --no-e2e > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err
python -c "
import json; d=json.load(open('gpurun_out/bench_$N.json')); print(d['value'], d['n_gpus'], 'sp', d.get('sp',{}).get('value'), d['clocks'])"; grep -v -i -E 'warn|OMP_NUM|\*\*\*|NCCL version|^$' gpurun_out/bench_$N.err | tail -5
