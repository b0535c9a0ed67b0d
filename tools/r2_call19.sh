#!/bin/bash
# rk_stage with cp.async row loads + exchange policy of HaloExchanger: tests, default bench (time step)
tag=${1:-r2_v29}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu -x > $out/pytest_gpu.txt 2>&1; tail -3 $out/pytest_gpu.txt
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 > $out/bench_air512.json 2> $out/bench_air512.err; python -c "import json;d=json.load(open('$out/bench_air512.json'));print(d['ms_per_step'],d['roofline']['kernels_ms'], d.get('time_step'))"
