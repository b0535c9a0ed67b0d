#!/bin/bash
# One measurement pass on the GPU box: tests, bench lines, launch list, full ncu capture of the staged kernels.
#   gpurun --timeout 1500 -- 'bash tools/measure_round.sh v11'
tag=${1:-vX}
out=gpurun_out/$tag
mkdir -p $out
python -m pytest tests -q -m gpu > $out/pytest_gpu.txt 2>&1; tail -2 $out/pytest_gpu.txt
python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_reference_air512.json 2> $out/bench_reference.err
python bench.py --steps 10 --warmup 3 > $out/bench_air512.json 2> $out/bench_air512.err
for w in air256 tg512_skew tg128_hybrid tg128_teno; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > $out/bench_$w.json 2>> $out/bench_other.err
done
python tools/parity_report.py > $out/parity_report.txt 2>&1
# launch list of the bench command (serialised, cold-cache: shares only)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches_air512.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > $out/ncu_launches.log 2>&1
# full capture of one launch of each staged kernel + the fused RK stage (256^3 per kernel replay keeps it short)
ncu --set full --clock-control none --import-source on -k regex:"euler_z_chem|euler_xy_bulk|diffusion_staged|rk_stage_rows" -c 5 \
    -o $out/staged_full python tools/exp_timestep_once.py > $out/ncu_full.log 2>&1
tail -2 $out/ncu_full.log
ls -la $out
