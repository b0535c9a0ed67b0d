#!/bin/bash
# Euler-z march: unroll factor of the plane loop (ring rotation by renaming instead of moves)
tag=${1:-r2_v27}
out=gpurun_out/$tag
mkdir -p $out
run() { timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 "$@" > $out/bench_$tagb.json 2> $out/bench_$tagb.err; python -c "import json;d=json.load(open('$out/bench_$tagb.json'));print('$tagb', d['ms_per_step'],d['roofline']['kernels_ms'])"; }
tagb=base; run
for v in mz2 mz3 mz6; do tagb=$v; run --variant $v; done
