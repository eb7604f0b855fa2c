#!/bin/bash
set -x
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > $O/final_tests.log 2>&1; tail -3 $O/final_tests.log
timeout 300 python __graft_entry__.py smoke > $O/final_smoke.log 2>&1; tail -3 $O/final_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/final_bench.log 2>&1; tail -c 400 $O/final_bench.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/final_gp_launches.csv python tools/gp_bench.py 40 16 bfloat16 2 > $O/final_gp_ncu.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'local_corr_mma' -c 2 -o $O/final_px python tools/kernel_bench.py --only gather --local-corr-path 2 --scales s4u --iters 1 > $O/final_px_ncu.log 2>&1
