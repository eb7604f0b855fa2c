#!/bin/bash
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > $O/final2_tests.log 2>&1; tail -3 $O/final2_tests.log
timeout 300 python __graft_entry__.py smoke > $O/final2_smoke.log 2>&1; tail -3 $O/final2_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/final2_bench.log 2>&1; tail -c 600 $O/final2_bench.log
