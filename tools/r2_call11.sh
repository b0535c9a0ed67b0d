#!/bin/bash
# cp.async diffusion kernels: L2 prefetch distance A/B
tag=${1:-r2_v17}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_parity_gpu.py -q -m gpu -x -k "staged_rhs_equals_per_task_sequence" > $out/pytest_staged.txt 2>&1; tail -4 $out/pytest_staged.txt
run() { timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 "$@" > $out/bench_$tagb.json 2> $out/bench_$tagb.err; python -c "import json;d=json.load(open('$out/bench_$tagb.json'));print('$tagb', d['ms_per_step'],d['roofline']['kernels_ms'])"; }
tagb=default; run
tagb=pf444; run --opt dprefetch=444
tagb=pf888; run --opt dprefetch=888
tagb=pf1776; run --opt dprefetch=1776
ls -la $out
