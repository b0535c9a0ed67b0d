#!/bin/bash
# One measurement pass of round 2 on the GPU box.   gpurun --timeout 1500 -- 'bash tools/r2_call.sh TAG [quick]'
tag=${1:-r2_vX}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/gpu.txt 2>&1
timeout 900 python -m pytest tests -q -m gpu -x > $out/pytest_gpu.txt 2>&1; tail -5 $out/pytest_gpu.txt
timeout 600 python bench.py --steps 10 --warmup 3 > $out/bench_air512.json 2> $out/bench_air512.err; tail -c 1500 $out/bench_air512.json
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --opt march=0 > $out/bench_air512_staged.json 2> $out/bench_air512_staged.err
timeout 300 python bench.py --workload air256_front --steps 10 --warmup 3 --no-cpu > $out/bench_air256_front.json 2> $out/bench_air256_front.err
timeout 300 python bench.py --workload air256_front --steps 10 --warmup 3 --no-cpu --opt march=0 > $out/bench_air256_front_staged.json 2>> $out/bench_air256_front.err
timeout 300 python bench.py --workload tg512_skew --steps 10 --warmup 3 --no-cpu > $out/bench_tg512_skew.json 2> $out/bench_tg512_skew.err
if [ "$2" != "quick" ]; then
timeout 200 python tools/exp_update_chemistry.py 32 > $out/update_chemistry_32.txt 2>&1
timeout 300 python tools/exp_rhs_once.py 256 2 > $out/rhs_once_256.txt 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"march_kernel" -c 2 -o $out/march_full \
    python tools/exp_rhs_once.py 256 1 > $out/ncu_march.log 2>&1
tail -2 $out/ncu_march.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"update_chemistry_kernel" -c 1 -o $out/chem_full \
    python tools/exp_update_chemistry.py 32 > $out/ncu_chem.log 2>&1
tail -2 $out/ncu_chem.log
fi
ls -la $out
