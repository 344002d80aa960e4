#!/bin/bash
N=8; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for V in 1 2 3; do
ND_NVLS_VARIANT=$V timeout 150 $TR --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --sync nvls --bucket-mb 256 --no-tokens 2> $OUT/sweep.err | grep '^{' > $OUT/sweep.json; python -c "
import json; d=json.load(open('$OUT/sweep.json')); print('nvls variant',$V,'ms',round(d['value'],3))"
done
for V in 1 2 3; do
ND_P2P8_VARIANT=$V timeout 150 $TR --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --sync p2p --no-tokens --no-e2e 2> $OUT/sweep.err | grep '^{' > $OUT/sweep.json; python -c "
import json; d=json.load(open('$OUT/sweep.json')); print('p2p8 variant',$V,'ms',round(d['value'],3))"
done
tail -3 $OUT/sweep.err
