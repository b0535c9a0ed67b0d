#!/bin/bash
tag=${1:-r2_v2}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests -q -m gpu -x > $out/pytest_gpu.txt 2>&1; tail -5 $out/pytest_gpu.txt
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu > $out/bench_air512.json 2> $out/bench_air512.err; python -c "import json;d=json.load(open('$out/bench_air512.json'));print(d['ms_per_step'],d['roofline']['kernels_ms'])"
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --variant mz3 > $out/bench_air512_mz3.json 2> $out/bench_air512_mz3.err; python -c "import json;d=json.load(open('$out/bench_air512_mz3.json'));print(d['ms_per_step'],d['roofline']['kernels_ms'])"
timeout 300 python tools/exp_rhs_once.py 256 2 > $out/rhs_once_256.txt 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"march_kernel" -c 2 -o $out/march_full \
    python tools/exp_rhs_once.py 256 1 > $out/ncu_march.log 2>&1
tail -2 $out/ncu_march.log
ls -la $out
