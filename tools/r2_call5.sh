#!/bin/bash
tag=${1:-r2_v5}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu -s > $out/pytest_gpu.txt 2>&1; grep -E "UpdateChemistry|passed|failed|FAILED" $out/pytest_gpu.txt | tail -12
for v in "" chem2 chem6; do timeout 300 python tools/exp_update_chemistry.py 96 $v > $out/update_chemistry_96_$v.txt 2>&1; echo "variant=$v $(cat $out/update_chemistry_96_$v.txt)"; done
ls -la $out
