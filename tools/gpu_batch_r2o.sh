#!/bin/bash
set -x
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > $O/r2o_tests.log 2>&1; tail -4 $O/r2o_tests.log
timeout 300 python tools/kernel_bench.py --only resample > $O/r2o_k3a.jsonl 2>&1; cat $O/r2o_k3a.jsonl | cut -c1-200
timeout 300 python __graft_entry__.py smoke > $O/r2o_smoke.log 2>&1; tail -3 $O/r2o_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/r2o_bench.log 2>&1; tail -c 800 $O/r2o_bench.log
