#!/bin/bash
# round 2, GPU run 3: the pruned library (8 tuning knobs, experiment build for the staged kernels, boundary options) through the
# whole GPU suite; where do the in-place and the A/B engine disagree at 8192x32768; better L2 probe
O=gpurun_out/r2_run3; mkdir -p $O
timeout 600 python tools/engine_diff.py dp 8192 32768 32 > $O/engine_diff_dp.log 2>&1; echo "diff dp rc=$?" | tee -a $O/rc.txt; tail -25 $O/engine_diff_dp.log
timeout 600 python tools/engine_diff.py dp 2048 8192 32 > $O/engine_diff_dp_small.log 2>&1; tail -12 $O/engine_diff_dp_small.log
timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -15 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt; tail -4 $O/smoke.log
timeout 900 python tools/l2_peak_probe.py > $O/l2_peak.log 2>&1; echo "l2 rc=$?" | tee -a $O/rc.txt; tail -3 $O/l2_peak.log; cp gpurun_out/l2_peak.json $O/ 2>/dev/null
