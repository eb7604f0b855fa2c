#!/bin/bash
set -x
O=gpurun_out
timeout 900 python -m pytest tests/test_local_corr_tile_gpu.py -m gpu -q > $O/r2e_tile_tests.log 2>&1; tail -5 $O/r2e_tile_tests.log
timeout 600 python -m pytest tests/test_gp_planes_gpu.py -m gpu -q -s > $O/r2e_gp_planes.log 2>&1; tail -8 $O/r2e_gp_planes.log
timeout 300 python tools/gp_bench.py 40 16 bfloat16 3 > $O/r2e_gp_bench.log 2>&1; cat $O/r2e_gp_bench.log
timeout 600 python -m pytest tests/test_kde_exact_gpu.py -m gpu -q > $O/r2e_kde_exact.log 2>&1; tail -3 $O/r2e_kde_exact.log
timeout 300 python tools/kernel_bench.py --only kde --kde-mode 0 > $O/r2e_kde_mode0.jsonl 2>&1; cat $O/r2e_kde_mode0.jsonl
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8,s4,s8u,s4u > $O/r2e_gather_path0.jsonl 2>&1; cat $O/r2e_gather_path0.jsonl
timeout 600 python tools/patch_diag.py > $O/r2e_patch_diag.log 2>&1; grep -v Warn $O/r2e_patch_diag.log | tail -14
timeout 1200 python -m pytest tests/test_refparity_gpu.py -m gpu -q -s > $O/r2e_refparity.log 2>&1; tail -4 $O/r2e_refparity.log
timeout 1200 python -m pytest tests/test_refparity_e2e_gpu.py -m gpu -q -s > $O/r2e_refparity_e2e.log 2>&1; tail -6 $O/r2e_refparity_e2e.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/r2e_bench.log 2>&1; tail -c 2500 $O/r2e_bench.log
