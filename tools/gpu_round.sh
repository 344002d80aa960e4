#!/bin/bash
# One GPU-box visit: smoke, GPU parity tests, bench (both arms), ncu launch list + one full capture.
# Usage (from the repo root, on the box):  bash tools/gpu_round.sh [tag]
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/gpu_$TAG.txt 2>&1
echo "== smoke" ; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke_$TAG.log
echo "== pytest -m gpu" ; timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest_$TAG.log
echo "== bench" ; timeout 900 python bench.py --steps 50 --warmup 5 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cat $OUT/bench_$TAG.json; tail -3 $OUT/bench_$TAG.err
echo "== bench reference arm" ; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "ref rc=$?"; cat $OUT/bench_ref_$TAG.json
echo "== ncu launch list"
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-tokens --e2e-steps 2"
timeout 600 $CMD > $OUT/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launches_$TAG.log 2>&1
echo "launch list rc=$?"
echo "== ncu full capture (150M-class model so that 40 replays stay short)"
CMD2="python bench.py --workload llama150m --steps 3 --warmup 3 --no-cpu-baseline --no-extras --no-tokens --e2e-steps 2"
timeout 600 $CMD2 > $OUT/plain150_$TAG.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 4 -c 2 -o $OUT/prof_$TAG -f $CMD2 > $OUT/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"
ls -la $OUT
