#!/bin/bash
# run 11: software-pipelined in-place kernels; e2e breakdown
mkdir -p gpurun_out
echo "=== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu.log
echo "=== pipe sweep"; timeout 900 python tools/pipe_sweep.py 2>&1 | tee gpurun_out/pipe_sweep.log | grep -E "^best|rror"
echo "=== bench"; timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print(d['value'], d['e2e'], d['sp']['value'], d['sp']['e2e'])"; tail -3 gpurun_out/bench.err
