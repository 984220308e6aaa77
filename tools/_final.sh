mkdir -p gpurun_out/r02y
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02y/pytest_gpu.log 2>&1; tail -2 gpurun_out/r02y/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/r02y/smoke.log 2>&1; tail -1 gpurun_out/r02y/smoke.log
timeout 300 python bench.py > gpurun_out/r02y/bench.json 2> gpurun_out/r02y/bench.err; cut -c1-330 gpurun_out/r02y/bench.json
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/r02y/bench20.json 2> gpurun_out/r02y/bench20.err; cut -c1-330 gpurun_out/r02y/bench20.json
timeout 250 python tools/round_cost.py > gpurun_out/r02y/round_cost.txt 2>&1
for n in 151552 606208 20; do echo "== samples (or log2) $n"; NF_PAIR_CFG=16,8,4,2 timeout 100 python tools/stamps.py $n 2>&1 | tail -7; done > gpurun_out/r02y/stamps.txt 2>&1
tail -3 gpurun_out/r02y/round_cost.txt
