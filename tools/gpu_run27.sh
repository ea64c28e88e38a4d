#!/bin/bash
# run 27: dataflow kernel, second version (release-add instead of fence + atomic, next item's counters read one item ahead)
mkdir -p gpurun_out
echo "=== pytest (dataflow)"; timeout 600 python -m pytest tests -m gpu -x -q -k "dataflow or every_kernel_variant" 2>&1 | tail -5 | tee gpurun_out/pytest_df.log
echo "=== sweep v2"; timeout 400 python tools/df_sweep.py v2 > gpurun_out/df_sweep_v2.log 2>&1; grep -E "best|default|Error|error" gpurun_out/df_sweep_v2.log | cut -c1-330
