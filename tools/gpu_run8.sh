#!/bin/bash
# run 8: in-place (AA) kernels -- parity first, then the comparison with A/B
mkdir -p gpurun_out
echo "=== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.log
echo "=== aa sweep"; timeout 900 python tools/aa_sweep.py 2>&1 | tee gpurun_out/aa_sweep.log | grep -E "^best|took|Error|error" 
