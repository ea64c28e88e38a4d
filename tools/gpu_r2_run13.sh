#!/bin/bash
# round 2, GPU run 13 (1 GPU): lbm_reynolds_column_static (no allocation, static device scratch): its test alone in a fresh process,
# then the whole GPU suite the way the round's end runs it (-x, the new file collected last)
O=gpurun_out/r2_run${RUN:-13}; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_no_alloc_alone.log 2>&1; echo "alone rc=$?" | tee -a $O/rc.txt; tail -15 $O/pytest_no_alloc_alone.log | cut -c1-250
timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a $O/rc.txt; tail -5 $O/pytest_gpu.log | cut -c1-250
