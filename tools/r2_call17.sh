#!/bin/bash
# HTR_QDIV (reciprocal-multiply in the production recovery / transport): tests, default workload, point kernels
tag=${1:-r2_v24}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu -x > $out/pytest_gpu.txt 2>&1; tail -3 $out/pytest_gpu.txt
run() { timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 "$@" > $out/bench_$tagb.json 2> $out/bench_$tagb.err; python -c "import json;d=json.load(open('$out/bench_$tagb.json'));print('$tagb', d['ms_per_step'],d['roofline']['kernels_ms'], d.get('time_step'))"; }
tagb=air512; run
tagb=air256; run --workload air256
timeout 300 python tools/exp_pointwise.py > $out/pointwise.txt 2>&1; tail -12 $out/pointwise.txt
timeout 300 python tools/parity_report.py > $out/parity_report.txt 2>&1; tail -25 $out/parity_report.txt
ls $out
