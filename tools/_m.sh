for lg in 20 21 24; do timeout 100 python tools/one.py default 32 $lg 200 2>&1 | tail -1; done
timeout 250 python tools/round_cost.py 2>&1 | tail -10
timeout 600 python -m pytest tests/test_gpu_pair.py tests/test_zz_no_alloc.py tests/test_gpu_elbo.py tests/test_gpu_rng.py -m gpu -x -q 2>&1 | tail -3
