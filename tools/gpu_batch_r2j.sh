#!/bin/bash
set -x
O=gpurun_out
timeout 900 python -m pytest tests/test_local_corr_tile_gpu.py tests/test_refparity_gpu.py tests/test_kernels_gpu.py -m gpu -q > $O/r2j_tests.log 2>&1; tail -4 $O/r2j_tests.log
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8,s4,s8u,s4u > $O/r2j_gather_path0.jsonl 2>&1; cat $O/r2j_gather_path0.jsonl | cut -c1-220
for d in 256 3840 7936; do timeout 300 python tools/kernel_bench.py --only gather --local-corr-path $d --scales s8u,s4u --iters 10 > $O/r2j_tile_dbg$d.jsonl 2>&1; grep '"flow": "coherent"' $O/r2j_tile_dbg$d.jsonl | cut -c1-200; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'local_corr_tile' -c 2 -o $O/r2j_tile_empty python tools/kernel_bench.py --only gather --local-corr-path 3840 --scales s4u --iters 1 > $O/r2j_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'local_corr_tile|local_corr_mma' -c 4 -o $O/r2j_tile_full python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s4u --iters 1 > $O/r2j_ncu2.log 2>&1
