#!/bin/bash
set -x
O=gpurun_out
timeout 600 python -m pytest tests/test_local_corr_tile_gpu.py tests/test_refparity_gpu.py tests/test_kernels_gpu.py -m gpu -q > $O/r2n_tests.log 2>&1; tail -4 $O/r2n_tests.log
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 0 --scales s16 > $O/r2n_gather_s16.jsonl 2>&1; cat $O/r2n_gather_s16.jsonl | cut -c1-220
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 2 --scales s16 > $O/r2n_gather_s16_p2.jsonl 2>&1; cat $O/r2n_gather_s16_p2.jsonl | cut -c1-220
timeout 300 python tools/kernel_bench.py --only resample,epilogue > $O/r2n_k3a.jsonl 2>&1; cat $O/r2n_k3a.jsonl | cut -c1-200
