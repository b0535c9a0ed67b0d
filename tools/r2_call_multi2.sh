#!/bin/bash
# multi-GPU check of the round's final build:  gpurun --gpus N --timeout 1200 -- 'bash tools/r2_call_multi2.sh TAG N'
tag=${1:-r2_multi}; N=${2:-2}
out=gpurun_out/$tag
mkdir -p $out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/check_multi_gpu_time_step.py > $out/time_step_parity_${N}gpu.txt 2>&1; tail -6 $out/time_step_parity_${N}gpu.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --e2e-steps 1 > $out/bench_air512_${N}gpu.json 2> $out/bench_${N}gpu.err; python -c "import json;d=json.load(open('$out/bench_air512_${N}gpu.json'));print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e'].get('ms_per_step'), d.get('time_step'), d.get('phases'))"
ls -la $out
