#!/bin/bash
# A/B of the cp.async diffusion staging (htr_set_option "dasync"): bit-identity test, bench lines, ncu of the new kernels
tag=${1:-r2_v14}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_parity_gpu.py -q -m gpu -x -k "staged_rhs_equals_per_task_sequence" > $out/pytest_staged.txt 2>&1; tail -4 $out/pytest_staged.txt
for o in dasync=0 dasync=1; do
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu --opt $o > $out/bench_air512_$o.json 2> $out/bench_$o.err; python -c "import json;d=json.load(open('$out/bench_air512_$o.json'));print('$o', d['ms_per_step'],d['roofline']['kernels_ms'])"
done
timeout 200 python tools/exp_rhs_once.py 256 3 dasync=1 > $out/rhs_once.txt 2>&1; tail -1 $out/rhs_once.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"diffusion_async" -c 2 \
    -o /tmp/dasync_full python tools/exp_rhs_once.py 256 1 dasync=1 > $out/ncu_full.log 2>&1
tail -2 $out/ncu_full.log
ncu -i /tmp/dasync_full.ncu-rep --page raw --csv > /tmp/dasync_raw.csv 2>/dev/null && python tools/ncu_summary.py /tmp/dasync_raw.csv --csv > $out/dasync_kernels_ncu_full_summary.csv
ncu -i /tmp/dasync_full.ncu-rep --page source --csv --print-source cuda,sass > /tmp/dasync_src.csv 2>/dev/null && python tools/ncu_source_hot.py /tmp/dasync_src.csv > $out/dasync_hot_source_lines.txt 2>&1
ncu -i /tmp/dasync_full.ncu-rep --page details --csv > $out/dasync_details.csv 2>/dev/null
ls -la $out
