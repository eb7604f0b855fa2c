#!/bin/bash
set -x
O=gpurun_out
for d in 0 256 512 1024 2048 3840; do timeout 300 python tools/kernel_bench.py --only gather --local-corr-path $d --scales s8u,s4u --iters 10 > $O/r2g_tile_dbg$d.jsonl 2>&1; grep coherent $O/r2g_tile_dbg$d.jsonl | grep -v incoh | cut -c1-200; done
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 > $O/r2g_gather_path0.jsonl 2>&1; cat $O/r2g_gather_path0.jsonl | cut -c1-220
timeout 300 python tools/gp_bench.py 40 16 bfloat16 2,3 > $O/r2g_gp_bench.log 2>&1; cat $O/r2g_gp_bench.log | cut -c1-250
timeout 300 python tools/gp_bench.py 48 8 bfloat16 2,3 > $O/r2g_gp_bench48.log 2>&1; cat $O/r2g_gp_bench48.log | cut -c1-250
timeout 1800 python -m pytest tests -m gpu -q -x > $O/r2g_tests.log 2>&1; tail -5 $O/r2g_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/r2g_bench.log 2>&1; tail -c 1500 $O/r2g_bench.log
