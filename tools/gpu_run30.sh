#!/bin/bash
# run 30: round-1 final verification on one GPU (parity suite, smoke, bench both arms)
mkdir -p gpurun_out
echo "=== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu.log
echo "=== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tee gpurun_out/smoke.log
echo "=== bench"; timeout 1500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print(d['value'], d['roofline']['frac'], d['e2e'], 'sp', d['sp']['value'], d['sp']['e2e']['value'], d['sp']['tuning'], d['ab_double_buffer']['dp']['value'], d['ab_double_buffer']['sp']['value']); print({k:v['value'] for k,v in d['weak_scaling_base_1gpu'].items()}); print(d['clocks'], d['cpu_baseline'])"; tail -3 gpurun_out/bench.err
echo "=== ref arm"; timeout 600 python bench.py --impl reference | tee gpurun_out/bench_ref.json | cut -c1-200
