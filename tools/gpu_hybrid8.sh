#!/bin/bash
N=8; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for F in 0.5 0.65 0.8; do
timeout 150 $TR --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --sync hybrid --nvls-fraction $F --bucket-mb 256 --no-tokens --no-extras 2> $OUT/hyb.err | grep '^{' > $OUT/bench_n8_hybrid_$F.json; python -c "
import json; d=json.load(open('$OUT/bench_n8_hybrid_$F.json')); print('hybrid f=$F ms', round(d['value'],3), 'frac', round(d['step_roofline']['frac_of_overlapped_bound'],3))"
done
grep -v "OMP_NUM\|^\*\*\*\|^$" $OUT/hyb.err | tail -3
