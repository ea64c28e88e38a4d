#!/bin/bash
# third GPU pass: PDL / CUDA-graph / persistent-v2 engines -- tests, sweep, bench
mkdir -p gpurun_out
echo "=== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
echo "=== pytest gpu"; timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -30 | tee gpurun_out/pytest_gpu.log
echo "=== sweep"; timeout 1500 python tools/sweep.py > gpurun_out/sweep.log 2>&1; tail -75 gpurun_out/sweep.log
echo "=== bench"; timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 3500 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
