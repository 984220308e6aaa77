mkdir -p gpurun_out/r02sp
for lg in 20 21 24; do timeout 100 python tools/one.py default 32 $lg 200 2>&1 | tail -1; done > gpurun_out/r02sp/final.txt 2>&1
cat gpurun_out/r02sp/final.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/r02sp/bench20.json 2>gpurun_out/r02sp/bench20.err; cut -c1-400 gpurun_out/r02sp/bench20.json
