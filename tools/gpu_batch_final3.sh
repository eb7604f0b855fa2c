#!/bin/bash
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > $O/final3_tests.log 2>&1; tail -2 $O/final3_tests.log
timeout 300 python __graft_entry__.py smoke > $O/final3_smoke.log 2>&1; tail -2 $O/final3_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/final3_bench.log 2>&1; tail -c 300 $O/final3_bench.log
timeout 600 python tools/kernel_bench.py --only gather --scales s8,s4,s8u,s4u > $O/final3_gather.jsonl 2>&1; cut -c1-210 $O/final3_gather.jsonl | tail -8
