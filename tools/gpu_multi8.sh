#!/bin/bash
# 8-GPU visit (8x charge: keep it short): bash tools/gpu_multi8.sh N tag
N=${1:-8}; TAG=${2:-r01}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > $OUT/topo_n${N}_$TAG.txt 2>&1
echo "== pytest sync (world $N only)"; timeout 300 python -m pytest tests/test_gpu_sync.py -x -q -k "\-$N]" 2>&1 | tail -4
echo "== bench p2p"; timeout 200 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --sync p2p 2> $OUT/bench_n${N}_p2p_$TAG.err | grep '^{' > $OUT/bench_n${N}_p2p_$TAG.json; cat $OUT/bench_n${N}_p2p_$TAG.json; grep -v "OMP_NUM\|^\*\*\*\|^$" $OUT/bench_n${N}_p2p_$TAG.err | tail -5
echo "== bench nccl bucket 256"; timeout 200 $TR --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --sync nccl --bucket-mb 256 --streams 3 2> $OUT/bench_n${N}_nccl256_$TAG.err | grep '^{' > $OUT/bench_n${N}_nccl256_$TAG.json; python -c "
import json
d=json.load(open('$OUT/bench_n${N}_nccl256_$TAG.json')); print('nccl ms',d['value'],'frac',d['step_roofline']['frac_of_overlapped_bound'])"
