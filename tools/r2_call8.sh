#!/bin/bash
tag=${1:-r2_v8}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu -x > $out/pytest_gpu.txt 2>&1; tail -4 $out/pytest_gpu.txt
run() { timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu "$@" > $out/bench_$tagb.json 2> $out/bench_$tagb.err; python -c "import json;d=json.load(open('$out/bench_$tagb.json'));print('$tagb', d['ms_per_step'],d['roofline']['kernels_ms'])"; }
tagb=diff3; run
tagb=diff3_off; run --opt diff3=0
tagb=diff3_d2; run --variant d2
ls $out
