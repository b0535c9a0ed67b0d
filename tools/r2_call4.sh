#!/bin/bash
tag=${1:-r2_v4}
out=gpurun_out/$tag
mkdir -p $out
timeout 1200 python -m pytest tests -q -m gpu -x -s > $out/pytest_gpu.txt 2>&1; tail -12 $out/pytest_gpu.txt
timeout 300 python tools/exp_update_chemistry.py 96 > $out/update_chemistry_96.txt 2>&1; cat $out/update_chemistry_96.txt
timeout 300 python tools/exp_update_chemistry.py 32 > $out/update_chemistry_32.txt 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"update_chemistry_kernel" -c 1 -o $out/chem_full \
    python tools/exp_update_chemistry.py 32 > $out/ncu_chem.log 2>&1
tail -2 $out/ncu_chem.log
ls -la $out
