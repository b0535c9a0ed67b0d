#!/bin/bash
# generic per-face kernels (TENO): register cap / residency A/B (build variants fb2 = 128 registers, fb3 = 80)
tag=${1:-r2_v19}
out=gpurun_out/$tag
mkdir -p $out
run() { timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 "$@" > $out/bench_$tagb.json 2> $out/bench_$tagb.err; python -c "import json;d=json.load(open('$out/bench_$tagb.json'));print('$tagb', d['ms_per_step'],d['roofline']['kernels_ms'])"; }
for v in "" fb2 fb3; do
  vv=${v:-base}; va=""; [ -n "$v" ] && va="--variant $v"
  tagb=tg128_teno_$vv; run --workload tg128_teno $va
  tagb=air256_front_$vv; run --workload air256_front $va
done
ls $out
