#!/bin/bash
set -x
O=gpurun_out
timeout 900 python -m pytest tests/test_local_corr_tile_gpu.py tests/test_refparity_gpu.py tests/test_gp_planes_gpu.py -m gpu -q > $O/r2k_tests.log 2>&1; tail -4 $O/r2k_tests.log
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8,s4,s8u,s4u > $O/r2k_gather_path0.jsonl 2>&1; cat $O/r2k_gather_path0.jsonl | cut -c1-220
timeout 300 python tools/kernel_bench.py --only gather --local-corr-path 256 --scales s8u,s4u --iters 10 > $O/r2k_tile_dbg256.jsonl 2>&1; grep '"flow": "coherent"' $O/r2k_tile_dbg256.jsonl | cut -c1-200
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'local_corr_tile|local_corr_mma' -c 4 -o $O/r2k_tile_full python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8u,s4u --iters 1 > $O/r2k_ncu2.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r2k_gp_launches.csv python tools/gp_bench.py 40 16 bfloat16 2 > $O/r2k_gp_ncu.log 2>&1
