#!/bin/bash
# cutDiv (reciprocal-multiply for the cut-off quotients of the TENO weights / sensors): tests, TENO + front + default workloads
tag=${1:-r2_v23}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu -x > $out/pytest_gpu.txt 2>&1; tail -3 $out/pytest_gpu.txt
run() { timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 "$@" > $out/bench_$tagb.json 2> $out/bench_$tagb.err; python -c "import json;d=json.load(open('$out/bench_$tagb.json'));print('$tagb', d['ms_per_step'],d['roofline']['kernels_ms'], d.get('time_step',{}).get('ms_per_time_step'))"; }
tagb=tg128_teno; run --workload tg128_teno
tagb=tg128_hybrid; run --workload tg128_hybrid
tagb=air256_front; run --workload air256_front
tagb=air512_front; run --workload air512_front
tagb=air512; run
timeout 300 python tools/exp_pointwise.py > $out/pointwise.txt 2>&1; tail -12 $out/pointwise.txt
ls $out
