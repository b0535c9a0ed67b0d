#!/bin/bash
# ncu of the per-face TENO kernels (128^3 ConstPropMix TENO) and of the listed pass on the front box
tag=${1:-r2_v22}
out=gpurun_out/$tag
mkdir -p $out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"flux_divergence_kernel" -c 3 \
    -o /tmp/teno_full python bench.py --workload tg128_teno --steps 1 --warmup 3 --no-cpu --e2e-steps 1 > $out/ncu_teno.log 2>&1
tail -2 $out/ncu_teno.log
ncu -i /tmp/teno_full.ncu-rep --page raw --csv > /tmp/teno_raw.csv 2>/dev/null && python tools/ncu_summary.py /tmp/teno_raw.csv --csv > $out/teno_kernels_ncu_full_summary.csv
python tools/ncu_summary.py /tmp/teno_raw.csv --grep "issue_stalled" | grep -E "^-----|per_issue_active.ratio" > $out/teno_kernels_stall_reasons.txt 2>&1
ncu -i /tmp/teno_full.ncu-rep --page source --csv --print-source cuda,sass > /tmp/teno_src.csv 2>/dev/null && python tools/ncu_source_hot.py /tmp/teno_src.csv 40 > $out/teno_hot_source_lines.txt 2>&1
ls -la $out
