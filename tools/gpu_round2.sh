#!/bin/bash
# One single-GPU box visit (round 2): smoke, GPU parity tests, bench (both arms), ncu launch list,
# full captures of the fused outer kernel and of the table-driven inner step.
# Usage (from the repo root, on the box):  bash tools/gpu_round2.sh [tag]
set -u
TAG=${1:-r02a}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/gpu_$TAG.txt 2>&1
free -g > $OUT/host_$TAG.txt; nproc >> $OUT/host_$TAG.txt
echo "== smoke" ; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke_$TAG.log
echo "== pytest -m gpu" ; timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest_$TAG.log
echo "== zz alone" ; timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
echo "== bench" ; timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cat $OUT/bench_$TAG.json; tail -3 $OUT/bench_$TAG.err
echo "== bench reference arm" ; timeout 800 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "ref rc=$?"; cat $OUT/bench_ref_$TAG.json
echo "== ncu launch list"
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-tokens --e2e-steps 2 --no-parity"
timeout 600 $CMD > $OUT/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launches_$TAG.log 2>&1
echo "launch list rc=$?"
echo "== ncu full capture: inner step (table-driven) on the 150M-class shapes"
CMD3="python tools/profile_inner.py"
timeout 300 $CMD3 > $OUT/plain_inner_$TAG.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"tile_table_kernel|table_kernel" -s 4 -c 2 -o $OUT/prof_inner_$TAG -f $CMD3 > $OUT/ncu_inner_$TAG.log 2>&1
echo "inner capture rc=$?"
ls -la $OUT | tail -20
