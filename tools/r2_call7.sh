#!/bin/bash
tag=${1:-r2_v7}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu -x > $out/pytest_gpu.txt 2>&1; tail -4 $out/pytest_gpu.txt
for o in tensor=1 tensor=0; do
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --opt $o > $out/bench_air512_$o.json 2> $out/bench_$o.err; python -c "import json;d=json.load(open('$out/bench_air512_$o.json'));print('$o', d['ms_per_step'],d['roofline']['kernels_ms'])"
done
timeout 300 python bench.py --workload air256_front --steps 10 --warmup 3 --no-cpu > $out/bench_air256_front.json 2>> $out/bench.err; python -c "import json;d=json.load(open('$out/bench_air256_front.json'));print('front', d['ms_per_step'],d['roofline']['kernels_ms'])"
ls $out
