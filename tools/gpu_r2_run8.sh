#!/bin/bash
# round 2, GPU run 8 (1 GPU): after removing the packed SP collide -- whole GPU suite, chain sweep for SP (scalar kernels) and for
# the L2-resident lattices, default-engine check at full size
O=gpurun_out/r2_run8; mkdir -p $O
timeout 3000 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -6 $O/pytest_gpu.log
timeout 900 python tools/chain_sweep.py sp all > $O/chain_sweep_sp.log 2>&1; echo "chain sweep sp rc=$?" | tee -a $O/rc.txt; cut -c1-200 $O/chain_sweep_sp.log
timeout 600 python tools/chain_sweep.py dp resident > $O/chain_sweep_dp_resident.log 2>&1; echo "chain sweep dp resident rc=$?" | tee -a $O/rc.txt; cut -c1-200 $O/chain_sweep_dp_resident.log
timeout 600 python tools/engine_diff.py sp 8192 32768 32 > $O/engine_diff_sp.log 2>&1; echo "diff sp rc=$?" | tee -a $O/rc.txt; tail -2 $O/engine_diff_sp.log | cut -c1-200
