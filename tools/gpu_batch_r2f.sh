#!/bin/bash
set -x
O=gpurun_out
timeout 600 python tools/config1_gp_paths.py > $O/r2f_gp_paths.log 2>&1; grep -v -i warn $O/r2f_gp_paths.log | tail -12
timeout 600 python -m pytest tests/test_local_corr_tile_gpu.py -m gpu -q > $O/r2f_tile_tests.log 2>&1; tail -3 $O/r2f_tile_tests.log
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_tiny.py -m gpu -q > $O/r2f_kernels.log 2>&1; tail -3 $O/r2f_kernels.log
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 2 --scales s2,s2u > $O/r2f_gather_v2.jsonl 2>&1; cat $O/r2f_gather_v2.jsonl
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 1 --scales s2,s2u > $O/r2f_gather_v1.jsonl 2>&1; cat $O/r2f_gather_v1.jsonl
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'local_corr_tile' -c 2 -o $O/r2f_tile python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8u --iters 1 > $O/r2f_ncu_tile.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:kde_exact_fused --launch-skip 8 -c 1 -o $O/r2f_kde_exact python tools/kernel_bench.py --only kde --kde-mode 0 --iters 4 > $O/r2f_ncu_kde.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2f_gp_launches.csv python tools/gp_bench.py 40 16 bfloat16 3 > $O/r2f_gp_ncu.log 2>&1
ls $O | grep r2f
