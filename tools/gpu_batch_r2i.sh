#!/bin/bash
set -x
O=gpurun_out
timeout 900 python -m pytest tests/test_local_corr_tile_gpu.py tests/test_refparity_gpu.py tests/test_gp_planes_gpu.py tests/test_kernels_gpu.py -m gpu -q > $O/r2i_tests.log 2>&1; tail -4 $O/r2i_tests.log
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8,s4,s8u,s4u > $O/r2i_gather_path0.jsonl 2>&1; cat $O/r2i_gather_path0.jsonl | cut -c1-220
for d in 256 3840; do timeout 300 python tools/kernel_bench.py --only gather --local-corr-path $d --scales s8u,s4u --iters 10 > $O/r2i_tile_dbg$d.jsonl 2>&1; grep '"flow": "coherent"' $O/r2i_tile_dbg$d.jsonl | cut -c1-200; done
timeout 300 python tools/gp_bench.py 40 16 bfloat16 2 > $O/r2i_gp_bench.log 2>&1; cat $O/r2i_gp_bench.log | cut -c1-250
