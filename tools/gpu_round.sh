#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, pipe-rate microbench, ncu launch list + full capture.
# Usage (from the repo root, under gpurun):  bash tools/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $OUT/nvsmi.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json
python bench.py --impl reference --steps 10 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; cat $OUT/bench_ref.json
[ -x tools/microbench ] && ./tools/microbench > $OUT/microbench.txt 2>&1
python bench.py --steps 20 --warmup 3 > $OUT/bench_short.json 2> $OUT/bench_short.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 20 --warmup 3 > $OUT/ncu_launches.log 2>&1
python bench.py --steps 20 --warmup 3 > $OUT/bench_short2.json 2> $OUT/bench_short2.err &&
ncu --set full --clock-control none --import-source on -k regex:elbo_shared_kernel -s 6 -c 2 -o $OUT/prof_elbo \
    python bench.py --steps 20 --warmup 3 > $OUT/ncu_full.log 2>&1
timeout 60 python tools/gemm_one.py 4096 1376 64 > $OUT/gemm_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tf32x3 -s 2 -c 1 -o $OUT/prof_gemm \
    python tools/gemm_one.py 4096 1376 64 > $OUT/ncu_gemm.log 2>&1
timeout 300 python tools/bench_configs.py > $OUT/configs.txt 2>&1
ls -la $OUT
