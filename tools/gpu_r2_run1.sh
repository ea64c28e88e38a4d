#!/bin/bash
# round 2, GPU run 1: the never-executed staged kernels (cp.async / TMA bulk), phase probe, L2 roof probe
mkdir -p gpurun_out/r2_run1; O=gpurun_out/r2_run1
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/gpu.txt
LBM_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k staged > $O/staged_test.log 2>&1; echo "staged test rc=$?" | tee -a $O/rc.txt
timeout 900 python tools/stage_phase_probe.py > $O/stage_phase.log 2>&1; echo "stage_phase rc=$?" | tee -a $O/rc.txt
timeout 600 python tools/l2_peak_probe.py > $O/l2_peak.log 2>&1; echo "l2 rc=$?" | tee -a $O/rc.txt
timeout 900 python tools/stage_probe.py > $O/stage_probe.log 2>&1; echo "stage_probe rc=$?" | tee -a $O/rc.txt
tail -5 $O/staged_test.log; tail -30 $O/stage_phase.log; tail -12 $O/l2_peak.log
