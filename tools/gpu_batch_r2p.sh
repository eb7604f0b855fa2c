#!/bin/bash
set -x
O=gpurun_out
timeout 600 python -m pytest tests/test_local_corr_tile_gpu.py tests/test_refparity_gpu.py -m gpu -q > $O/r2p_tests.log 2>&1; tail -3 $O/r2p_tests.log
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s8,s4,s8u,s4u > $O/r2p_gather_path0.jsonl 2>&1; cat $O/r2p_gather_path0.jsonl | cut -c1-220
