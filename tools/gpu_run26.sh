#!/bin/bash
# run 26: dataflow kernel (one launch per macro-step): parity, then sweeps against the launch-per-level engine
mkdir -p gpurun_out
echo "=== pytest (dataflow first)"; timeout 600 python -m pytest tests -m gpu -x -q -k "dataflow or every_kernel_variant" 2>&1 | tail -15 | tee gpurun_out/pytest_df.log
echo "=== sweep small"; timeout 400 python tools/df_sweep.py small > gpurun_out/df_sweep_small.log 2>&1; grep -E "best|default|Error|error" gpurun_out/df_sweep_small.log | cut -c1-330
echo "=== sweep large"; timeout 500 python tools/df_sweep.py large > gpurun_out/df_sweep_large.log 2>&1; grep -E "best|default|Error|error" gpurun_out/df_sweep_large.log | cut -c1-330
echo "=== pytest -m gpu (all)"; timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/pytest_gpu.log
