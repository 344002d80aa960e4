#!/bin/bash
# N-GPU visit of round 2: bash tools/gpu_multi_r2.sh N tag [what...]   (what: tests bench sweep nccl micro cfg)
N=${1:-2}; TAG=${2:-r02}; shift 2; WHAT=${*:-tests bench sweep}
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
FAST="--steps 20 --warmup 5 --no-tokens --no-extras"
free -g | head -2 > $OUT/host_n${N}_$TAG.txt; nproc >> $OUT/host_n${N}_$TAG.txt
run_bench() {  # name port args...
  local name=$1 port=$2; shift 2
  timeout 400 $TR --master-port $port bench.py --gpus $N "$@" 2> $OUT/bench_n${N}_${name}_$TAG.err | grep '^{' > $OUT/bench_n${N}_${name}_$TAG.json
  python - <<PY
import json
try:
    d=json.load(open('$OUT/bench_n${N}_${name}_$TAG.json'))
    r=d['roofline']; p=d.get('parity_check') or {}; nv=d.get('nvlink_measured') or {}; e=d.get('e2e') or {}
    print('$name', 'ms=%.3f'%d['value'], 'frac=%.3f'%r['frac'], 'parity=%s ulp=%s'%(p.get('ok'),p.get('max_ulp')), 'traffic/alg=%s'%nv.get('traffic_over_algorithmic'), 'e2e=%s ceil=%s'%(e.get('value'), e.get('host_link_ceiling_ms')), p.get('error',''), nv.get('error',''))
except Exception as ex:
    print('$name', 'FAILED', ex)
PY
  grep -v "OMP_NUM\|^\*\*\*\|^$" $OUT/bench_n${N}_${name}_$TAG.err | tail -4
}
for w in $WHAT; do case $w in
tests)
  echo "== pytest sync (world $N)"
  ND_SYNC_WORLDS=$N ND_SYNC_TEST_LOG=$OUT/sync_tests_n%W_$TAG.json timeout 1400 python -m pytest tests/test_gpu_sync.py -q > $OUT/pytest_sync_n${N}_$TAG.log 2>&1
  echo "pytest rc=$?"; tail -12 $OUT/pytest_sync_n${N}_$TAG.log; cat $OUT/sync_tests_n${N}_$TAG.json 2>/dev/null | tr -d '\n' | cut -c1-3000; echo;;
bench)
  run_bench pull 29511 $FAST --sync p2p --p2p-exchange pull
  run_bench push 29512 $FAST --sync p2p --p2p-exchange push;;
sweep)
  for v in 1 2 3; do ND_PUSH_VARIANT=$v run_bench pushv$v 2952$v $FAST --sync p2p --p2p-exchange push --no-e2e --no-parity; done;;
nccl)
  run_bench nccl64 29531 $FAST --sync nccl --bucket-mb 64 --no-e2e
  run_bench nccl256 29532 $FAST --sync nccl --bucket-mb 256 --streams 3 --no-e2e;;
hybrid)
  run_bench hybrid 29541 $FAST --sync hybrid --nvls-fraction 0.5 --no-e2e
  run_bench nvls 29542 $FAST --sync nvls --no-e2e;;
cfg)
  run_bench bf16 29551 $FAST --sync p2p --dtype bf16 --no-e2e
  run_bench cfg2_150m_bf16 29552 $FAST --sync p2p --dtype bf16 --workload llama150m --no-e2e
  [ $N -ge 4 ] && run_bench cfg5_rep2 29553 $FAST --sync p2p --replicas 2 --no-e2e;;
push)
  run_bench push 29512 --steps 10 --warmup 3 --no-tokens --no-extras --no-e2e --sync p2p --p2p-exchange push;;
nccl256)
  run_bench nccl256 29532 $FAST --sync nccl --bucket-mb 256 --streams 3 --no-e2e;;
hybrid65)
  run_bench hybrid65 29543 $FAST --sync hybrid --nvls-fraction 0.65 --no-e2e;;
default)
  timeout 600 $TR --master-port 29571 bench.py --gpus $N --steps 20 --warmup 5 2> $OUT/bench_n${N}_default_$TAG.err | grep '^{' > $OUT/bench_n${N}_default_$TAG.json; cut -c1-300 $OUT/bench_n${N}_default_$TAG.json; echo
  python - <<PY
import json
d=json.load(open('$OUT/bench_n${N}_default_$TAG.json')); p=d.get('parity_check') or {}; e=d.get('e2e') or {}; t=d.get('tokens_per_s') or {}
print('default ms=%.3f frac=%.3f parity=%s e2e=%s ceil=%s tokens=%s wait_ms=%s sync_ms=%s' % (d['value'], d['roofline']['frac'], p.get('ok'), e.get('value'), e.get('host_link_ceiling_ms'), t.get('value'), t.get('avg_sync_wait_ms_rank0'), t.get('avg_sync_ms')), t.get('error',''))
PY
  ;;
ref3)
  timeout 600 $TR --master-port 29572 bench.py --impl reference --gpus $N --steps 3 --warmup 1 2> $OUT/bench_n${N}_ref3_$TAG.err | grep '^{' > $OUT/bench_n${N}_ref3_$TAG.json; cut -c1-2000 $OUT/bench_n${N}_ref3_$TAG.json; tail -3 $OUT/bench_n${N}_ref3_$TAG.err;;
fold)
  run_bench fold 29581 $FAST --sync p2p --no-e2e
  run_bench nofold 29582 $FAST --sync p2p --no-e2e --p2p-no-fold
  run_bench fold_150m 29583 $FAST --sync p2p --no-e2e --workload llama150m
  run_bench nofold_150m 29584 $FAST --sync p2p --no-e2e --workload llama150m --p2p-no-fold
  run_bench fold_tiny 29585 $FAST --sync p2p --no-e2e --workload tiny
  run_bench nofold_tiny 29586 $FAST --sync p2p --no-e2e --workload tiny --p2p-no-fold;;
micro)
  timeout 600 $TR --master-port 29561 tools/microbench.py > $OUT/microbench_n${N}_$TAG.log 2>&1; tail -30 $OUT/microbench_n${N}_$TAG.log;;
full)
  timeout 600 $TR --master-port 29571 bench.py --gpus $N --steps 20 --warmup 5 2> $OUT/bench_n${N}_default_$TAG.err | grep '^{' > $OUT/bench_n${N}_default_$TAG.json; cut -c1-600 $OUT/bench_n${N}_default_$TAG.json
  timeout 860 $TR --master-port 29572 bench.py --impl reference --gpus $N --steps 20 --warmup 5 2> $OUT/bench_n${N}_ref_$TAG.err | grep '^{' > $OUT/bench_n${N}_ref_$TAG.json; cut -c1-400 $OUT/bench_n${N}_ref_$TAG.json;;
esac; done
