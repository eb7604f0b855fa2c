#!/bin/bash
# one gpurun call: tile-kernel correctness, patch diagnosis, parity tests, per-kernel timings, ncu captures, bench
set -x
O=gpurun_out
timeout 900 python -m pytest tests/test_local_corr_tile_gpu.py -m gpu -q -x > $O/r2d_tile_tests.log 2>&1; tail -5 $O/r2d_tile_tests.log
timeout 600 python tools/patch_diag.py > $O/r2d_patch_diag.log 2>&1; grep -v Warn $O/r2d_patch_diag.log | tail -14
timeout 1200 python -m pytest tests/test_refparity_gpu.py -m gpu -q -s > $O/r2d_refparity.log 2>&1; tail -4 $O/r2d_refparity.log
timeout 900 python -m pytest tests/test_refparity_e2e_gpu.py -m gpu -q -s > $O/r2d_refparity_e2e.log 2>&1; tail -4 $O/r2d_refparity_e2e.log
timeout 300 python __graft_entry__.py smoke > $O/r2d_smoke.log 2>&1; tail -3 $O/r2d_smoke.log
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 2 > $O/r2d_gather_path2.jsonl 2>&1
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8,s4,s8u,s4u > $O/r2d_gather_path0.jsonl 2>&1
for m in 0 3 4; do timeout 300 python tools/kernel_bench.py --only kde --kde-mode $m > $O/r2d_kde_mode$m.jsonl 2>&1; done
cat $O/r2d_gather_path2.jsonl $O/r2d_gather_path0.jsonl $O/r2d_kde_mode*.jsonl
timeout 900 python bench.py --steps 10 --warmup 3 > $O/r2d_bench.log 2>&1; tail -c 3000 $O/r2d_bench.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:kde_exact_fused -c 2 -o $O/r2d_kde_exact python tools/kernel_bench.py --only kde --kde-mode 0 --iters 4 > $O/r2d_ncu_kde.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'local_corr_tile|gather_only|gather_rowstage' -c 12 -o $O/r2d_k2 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8u,s2u,s1u --iters 1 > $O/r2d_ncu_k2.log 2>&1
ls -la $O
