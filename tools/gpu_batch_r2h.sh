#!/bin/bash
set -x
O=gpurun_out
timeout 600 python -m pytest tests/test_local_corr_tile_gpu.py tests/test_refparity_gpu.py -m gpu -q > $O/r2h_tests.log 2>&1; tail -3 $O/r2h_tests.log
for d in 0 3840 7936 16128; do timeout 300 python tools/kernel_bench.py --only gather --local-corr-path $d --scales s8u,s4u --iters 10 > $O/r2h_tile_dbg$d.jsonl 2>&1; grep '"flow": "coherent"' $O/r2h_tile_dbg$d.jsonl | cut -c1-200; done
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8,s4,s8u,s4u,s1,s1u > $O/r2h_gather_path0.jsonl 2>&1; cat $O/r2h_gather_path0.jsonl | cut -c1-220
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 2 --scales s8u,s4u > $O/r2h_gather_path2.jsonl 2>&1; cat $O/r2h_gather_path2.jsonl | cut -c1-220
