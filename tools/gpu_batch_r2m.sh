#!/bin/bash
set -x
O=gpurun_out
timeout 300 python -m pytest tests/test_refparity_gpu.py tests/test_kernels_gpu.py -m gpu -q -k "resample or epilogue or update" > $O/r2m_tests.log 2>&1; tail -3 $O/r2m_tests.log
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras > $O/r2m_bench_plain.log 2>&1; tail -c 300 $O/r2m_bench_plain.log
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file $O/r2m_bench_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras > $O/r2m_bench_ncu.log 2>&1
ls -la $O/r2m_bench_launches.csv
