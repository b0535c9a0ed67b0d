#!/bin/bash
tag=${1:-r2_v6}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu -s > $out/pytest_gpu.txt 2>&1; grep -E "Sod|UpdateChemistry (prod|strict)|passed|failed|FAILED|Error" $out/pytest_gpu.txt | cut -c1-300 | tail -14
ls -la $out
