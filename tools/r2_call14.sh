#!/bin/bash
# lane-cooperative TENO kernels: parity (coop vs per-face vs oracle), then the TENO / front workloads with coop = 1 / 0
tag=${1:-r2_v21}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_coop_gpu.py -q -m gpu > $out/pytest_coop.txt 2>&1; tail -15 $out/pytest_coop.txt
timeout 1500 python -m pytest tests -q -m gpu -x > $out/pytest_gpu.txt 2>&1; tail -3 $out/pytest_gpu.txt
run() { timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 1 "$@" > $out/bench_$tagb.json 2> $out/bench_$tagb.err; python -c "import json;d=json.load(open('$out/bench_$tagb.json'));print('$tagb', d['ms_per_step'],d['roofline']['kernels_ms'])"; }
for c in 1 0; do
  tagb=tg128_teno_coop$c; run --workload tg128_teno --opt coop=$c
  tagb=air256_front_coop$c; run --workload air256_front --opt coop=$c
  tagb=air512_front_coop$c; run --workload air512_front --opt coop=$c
done
ls $out
