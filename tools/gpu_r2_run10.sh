#!/bin/bash
# round 2, GPU run 10 (1 GPU): per-level seam-chain events on the last column piece -- whole GPU suite, chain sweep (small lattices are
# where the chain was exposed), engine agreement at full size
O=gpurun_out/r2_run10; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -6 $O/pytest_gpu.log
timeout 900 python tools/chain_sweep.py both all > $O/chain_sweep.log 2>&1; echo "chain sweep rc=$?" | tee -a $O/rc.txt; grep -E '"chains": [12], "H": 0' $O/chain_sweep.log | cut -c1-200; tail -1 $O/chain_sweep.log
timeout 600 python tools/engine_diff.py dp 8192 32768 32 > $O/engine_diff_dp.log 2>&1; echo "diff dp rc=$?" | tee -a $O/rc.txt; tail -1 $O/engine_diff_dp.log
timeout 600 python tools/engine_diff.py sp 8192 32768 32 > $O/engine_diff_sp.log 2>&1; echo "diff sp rc=$?" | tee -a $O/rc.txt; tail -1 $O/engine_diff_sp.log
