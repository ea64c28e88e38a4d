#!/bin/bash
# round 2, GPU run 7 (1 GPU): concurrent tile chains inside one context (lbm_tuning.tb_chains) -- sweep with checksums, the SP
# packed FP32x2 collide against the scalar build, the whole GPU suite with both
O=gpurun_out/r2_run7; mkdir -p $O
timeout 1200 python tools/chain_sweep.py both all > $O/chain_sweep.log 2>&1; echo "chain sweep rc=$?" | tee -a $O/rc.txt; cut -c1-200 $O/chain_sweep.log; cp gpurun_out/chain_sweep.json $O/ 2>/dev/null
timeout 900 python tools/sp_pack_probe.py > $O/sp_pack.log 2>&1; echo "sp pack rc=$?" | tee -a $O/rc.txt; cut -c1-200 $O/sp_pack.log
timeout 600 python tools/engine_diff.py dp 8192 32768 32 > $O/engine_diff_dp.log 2>&1; echo "diff dp rc=$?" | tee -a $O/rc.txt; tail -2 $O/engine_diff_dp.log | cut -c1-200
timeout 600 python tools/engine_diff.py sp 8192 32768 32 > $O/engine_diff_sp.log 2>&1; echo "diff sp rc=$?" | tee -a $O/rc.txt; tail -2 $O/engine_diff_sp.log | cut -c1-200
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -12 $O/pytest_gpu.log
