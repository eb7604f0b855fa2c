#!/bin/bash
set -x
O=gpurun_out
timeout 600 python -m pytest tests/test_gp_planes_gpu.py tests/test_kernels_gpu.py tests/test_fullsize_gpu.py tests/test_local_corr_tile_gpu.py -m gpu -q > $O/r2l_tests.log 2>&1; tail -4 $O/r2l_tests.log
timeout 300 python tools/gp_bench.py 40 16 bfloat16 2,3 > $O/r2l_gp_bench.log 2>&1; cat $O/r2l_gp_bench.log | cut -c1-250
timeout 300 python tools/gp_bench.py 48 8 bfloat16 2 > $O/r2l_gp_bench48.log 2>&1; cat $O/r2l_gp_bench48.log | cut -c1-250
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8u,s4u > $O/r2l_gather_path0.jsonl 2>&1; cat $O/r2l_gather_path0.jsonl | cut -c1-220
timeout 900 python bench.py --steps 10 --warmup 3 > $O/r2l_bench.log 2>&1; tail -c 1200 $O/r2l_bench.log
