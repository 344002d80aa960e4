#!/bin/bash
N=8; TAG=${1:-r01b}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== pytest sync (world 8 subset)"; timeout 300 python -m pytest tests/test_gpu_sync.py -x -q -k "p2p-64.0-8 or nvls-64.0-8 or nccl-64.0-8" 2>&1 | tail -4
for B in 256 64; do
echo "== bench nvls bucket $B"; timeout 200 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --sync nvls --bucket-mb $B --no-tokens 2> $OUT/bench_n${N}_nvls${B}_$TAG.err | grep '^{' > $OUT/bench_n${N}_nvls${B}_$TAG.json; python -c "
import json; d=json.load(open('$OUT/bench_n${N}_nvls${B}_$TAG.json')); print('nvls bucket',$B,'ms',d['value'],'frac',d['step_roofline']['frac_of_overlapped_bound'])"; grep -v "OMP_NUM\|^\*\*\*\|^$" $OUT/bench_n${N}_nvls${B}_$TAG.err | tail -4
done
echo "== default bench (p2p)"; timeout 400 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 2> $OUT/bench_n${N}_default_$TAG.err | grep '^{' > $OUT/bench_n${N}_default_$TAG.json; python -c "
import json; d=json.load(open('$OUT/bench_n${N}_default_$TAG.json')); print('p2p ms',d['value'],'e2e',d['e2e']['value'],'tokens/s',d['tokens_per_s'].get('value'), d['tokens_per_s'].get('outer_sync_ms'), d['tokens_per_s'].get('error'))"; grep -v "OMP_NUM\|^\*\*\*\|^$\|Warn\|warn" $OUT/bench_n${N}_default_$TAG.err | tail -4
