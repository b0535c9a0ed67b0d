#!/bin/bash
tag=${1:-r2_v3}
out=gpurun_out/$tag
mkdir -p $out
timeout 1200 python -m pytest tests -q -m gpu -x -s > $out/pytest_gpu.txt 2>&1; tail -8 $out/pytest_gpu.txt
for m in 1 5; do
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --opt march=$m > $out/bench_air512_march$m.json 2> $out/bench_air512_march$m.err; python -c "import json;d=json.load(open('$out/bench_air512_march$m.json'));print($m, d['ms_per_step'],d['roofline']['kernels_ms'])"
done
ls -la $out
