#!/bin/bash
set -x
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > $O/r2u_tests.log 2>&1; tail -3 $O/r2u_tests.log
timeout 300 python __graft_entry__.py smoke > $O/r2u_smoke.log 2>&1; tail -3 $O/r2u_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/r2u_bench.log 2>&1; tail -c 600 $O/r2u_bench.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2u_bench_ref.log 2>&1; tail -c 600 $O/r2u_bench_ref.log
