#!/bin/bash
# Multi-GPU visit: bash tools/gpu_multi.sh N tag
N=${1:-2}; TAG=${2:-r01}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== pytest sync"; timeout 600 python -m pytest tests/test_gpu_sync.py -x -q 2>&1 | tail -8
echo "== bench p2p"; timeout 300 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --sync p2p > $OUT/bench_n${N}_p2p_$TAG.json 2> $OUT/bench_n${N}_p2p_$TAG.err; echo rc=$?; cat $OUT/bench_n${N}_p2p_$TAG.json; grep -v "OMP_NUM\|^\*\*\*\|^$" $OUT/bench_n${N}_p2p_$TAG.err | tail -5
for B in 64 256 1024; do
echo "== bench nccl bucket $B"; timeout 300 $TR --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --sync nccl --bucket-mb $B --streams 3 > $OUT/bench_n${N}_nccl${B}_$TAG.json 2> $OUT/bench_n${N}_nccl${B}_$TAG.err; echo rc=$?; python -c "
import json,sys
d=json.load(open('$OUT/bench_n${N}_nccl${B}_$TAG.json')); print('nccl bucket',$B,'ms',d['value'],'frac',d['step_roofline']['frac_of_overlapped_bound'])"
done
