#!/bin/bash
# One GPU-box visit for the round-2 build: parity tests, smoke, both bench arms, ncu launch list + full capture of
# the headline kernel, phase stamps.  Usage (from the repo root, under gpurun):  bash tools/gpu_round2.sh [tag]
TAG=${1:-r02z}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $OUT/nvsmi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log
timeout 300 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json
timeout 300 python bench.py --impl reference --steps 10 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; cat $OUT/bench_ref.json
timeout 300 python bench.py --steps 20 --warmup 3 > $OUT/bench_short.json 2> $OUT/bench_short.err &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 20 --warmup 3 > $OUT/ncu_launches.log 2>&1
cat $OUT/bench_short.json
timeout 300 python bench.py --steps 20 --warmup 3 > $OUT/bench_short2.json 2> $OUT/bench_short2.err &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:elbo_pair_kernel -s 6 -c 2 -o $OUT/prof_pair \
    python bench.py --steps 20 --warmup 3 > $OUT/ncu_full.log 2>&1
[ -f tools/stamps.py ] && NF_PAIR_CFG=16,8,4,2 timeout 120 python tools/stamps.py > $OUT/stamps.txt 2>&1
[ -f tools/sweep.py ] && timeout 300 python tools/sweep.py > $OUT/sweep.txt 2>&1
timeout 300 python tools/bench_configs.py > $OUT/configs.txt 2>&1
ls -la $OUT
