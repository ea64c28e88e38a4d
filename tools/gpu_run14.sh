#!/bin/bash
# run 14: 2-D temporal blocking (column strips) -- parity, then the wide lattices
mkdir -p gpurun_out
echo "=== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu.log
echo "=== tb2d sweep"; timeout 1500 python tools/tb2d_sweep.py both 2>&1 | tee gpurun_out/tb2d_sweep.log | grep -E "^best|rror"
