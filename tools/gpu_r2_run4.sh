#!/bin/bash
# round 2, GPU run 4: after the lbm_create memset-ordering fix -- engines must agree at 8192x32768; the whole GPU suite; L2 probe
# with a persistent mode; tile sweep (strip width x budget x depth x evict-first stores)
O=gpurun_out/r2_run4; mkdir -p $O
timeout 600 python tools/engine_diff.py dp 8192 32768 32 > $O/engine_diff_dp.log 2>&1; echo "diff dp rc=$?" | tee -a $O/rc.txt; tail -8 $O/engine_diff_dp.log | cut -c1-200
timeout 600 python tools/engine_diff.py sp 8192 32768 32 > $O/engine_diff_sp.log 2>&1; echo "diff sp rc=$?" | tee -a $O/rc.txt; tail -3 $O/engine_diff_sp.log | cut -c1-200
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -15 $O/pytest_gpu.log
timeout 900 python tools/l2_peak_probe.py > $O/l2_peak.log 2>&1; echo "l2 rc=$?" | tee -a $O/rc.txt; tail -2 $O/l2_peak.log; cp gpurun_out/l2_peak.json $O/ 2>/dev/null
timeout 1500 python tools/tile_sweep.py both > $O/tile_sweep.log 2>&1; echo "tile rc=$?" | tee -a $O/rc.txt; grep best $O/tile_sweep.log; cp gpurun_out/tile_sweep.json $O/ 2>/dev/null
