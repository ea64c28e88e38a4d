#!/bin/bash
# run 9: temporal blocking in L2 (wavefront schedule over the in-place kernels)
mkdir -p gpurun_out
echo "=== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.log
echo "=== tb sweep"; timeout 1200 python tools/tb_sweep.py all 2>&1 | tee gpurun_out/tb_sweep.log | grep -E "^best|took|rror"
