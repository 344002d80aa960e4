#!/bin/bash
N=8; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 150 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --no-tokens --no-extras --no-e2e "$@" 2> $OUT/$name.err | grep '^{' > $OUT/$name.json; python -c "
import json; d=json.load(open('$OUT/$name.json')); print('$name', 'ms', round(d['value'],3), 'frac', round(d['step_roofline']['frac_of_overlapped_bound'],3), 'nvlink GB/s', round(d['step_roofline']['nvlink_algorithmic_gbs_per_dir']), d['config']['parallelism'])"; grep -v "OMP_NUM\|^\*\*\*\|^$\|Warn\|warn" $OUT/$name.err | tail -2; }
run bench_n8_cfg2_llama150m_bf16_r01 --workload llama150m --dtype bf16
