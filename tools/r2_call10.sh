#!/bin/bash
# cp.async diffusion kernels: planes per CTA, 12-row tiles; stall reasons of the staged kernels
tag=${1:-r2_v16}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_parity_gpu.py -q -m gpu -x -k "staged_rhs_equals_per_task_sequence" > $out/pytest_staged.txt 2>&1; tail -4 $out/pytest_staged.txt
run() { timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu "$@" > $out/bench_$tagb.json 2> $out/bench_$tagb.err; python -c "import json;d=json.load(open('$out/bench_$tagb.json'));print('$tagb', d['ms_per_step'],d['roofline']['kernels_ms'])"; }
tagb=default; run
tagb=dplanes2; run --opt dplanes=2
tagb=dplanes4; run --opt dplanes=4
tagb=dplanes8; run --opt dplanes=8
tagb=tj12; run --variant tj12
tagb=tj12_dplanes4; run --variant tj12 --opt dplanes=4
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"euler_z_march|euler_xy_bulk|diffusion_async" -c 4 \
    -o /tmp/staged_full python tools/exp_rhs_once.py 256 1 > $out/ncu_full.log 2>&1
tail -2 $out/ncu_full.log
ncu -i /tmp/staged_full.ncu-rep --page raw --csv > /tmp/staged_raw.csv 2>/dev/null && python tools/ncu_summary.py /tmp/staged_raw.csv --csv > $out/staged_kernels_ncu_full_summary.csv
python tools/ncu_summary.py /tmp/staged_raw.csv --grep "issue_stalled,l1tex__,smsp__inst_executed_pipe,sm__inst_executed_pipe,lsu,pipe_" > $out/staged_kernels_stalls.txt 2>&1
ncu -i /tmp/staged_full.ncu-rep --page source --csv --print-source cuda,sass > /tmp/staged_src.csv 2>/dev/null && python tools/ncu_source_hot.py /tmp/staged_src.csv 40 > $out/hot_source_lines.txt 2>&1
ls -la $out
