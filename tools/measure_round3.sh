#!/bin/bash
# Round-2 final measurement pass on the GPU box: tests, bench lines (both arms), launch list, full ncu capture.
#   gpurun --timeout 2400 -- 'bash tools/measure_round3.sh r2_v18'
tag=${1:-r2_vX}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu > $out/pytest_gpu.txt 2>&1; tail -3 $out/pytest_gpu.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > $out/pytest_no_alloc_alone.txt 2>&1; tail -1 $out/pytest_no_alloc_alone.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_reference_air512.json 2> $out/bench_reference.err
timeout 900 python bench.py --steps 10 --warmup 3 > $out/bench_air512.json 2> $out/bench_air512.err
python -c "import json;d=json.load(open('$out/bench_air512.json'));print(d['ms_per_step'],d['value'],d['e2e'],d['roofline'].get('frac'),d.get('parity'),d.get('time_step'))"
timeout 300 python tools/exp_update_chemistry.py 96 > $out/update_chemistry_timing.txt 2>&1; tail -1 $out/update_chemistry_timing.txt
timeout 300 python tools/exp_pointwise.py > $out/pointwise.txt 2>&1
for w in air256 air256_front air512_front air512_tanh tg512_skew tg128_hybrid tg128_teno; do
  timeout 400 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > $out/bench_$w.json 2>> $out/bench_other.err
  python -c "import json;d=json.load(open('$out/bench_$w.json'));print('$w', d['ms_per_step'],d['roofline']['kernels_ms'])"
done
timeout 600 python tools/parity_report.py > $out/parity_report.txt 2>&1; tail -3 $out/parity_report.txt
# launch list of the bench command (serialised, cold-cache: shares only)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches_air512.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > $out/ncu_launches.log 2>&1
# full capture of one launch of each staged kernel + the fused RK stage on the bench box; exported to CSV on the box
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"euler_z_march|euler_xy_bulk|diffusion_async|rk_stage_rows" -c 5 \
    -o /tmp/staged_full python tools/exp_timestep_once.py > $out/ncu_full.log 2>&1
tail -2 $out/ncu_full.log
ncu -i /tmp/staged_full.ncu-rep --page raw --csv > /tmp/staged_raw.csv 2>/dev/null && python tools/ncu_summary.py /tmp/staged_raw.csv --csv > $out/staged_kernels_ncu_full_summary.csv
python tools/ncu_summary.py /tmp/staged_raw.csv --grep "issue_stalled" | grep -E "^-----|per_issue_active.ratio" > $out/staged_kernels_stall_reasons.txt 2>&1
ncu -i /tmp/staged_full.ncu-rep --page source --csv --print-source cuda,sass > /tmp/staged_src.csv 2>/dev/null && python tools/ncu_source_hot.py /tmp/staged_src.csv > $out/hot_source_lines.txt 2>&1
du -sh $out; ls $out
