#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests/test_local_corr_tile_gpu.py tests/test_kernels_gpu.py tests/test_refparity_gpu.py -q -x -k "corr or gather or refiner or k2 or K2" > $O/r3r_tests.log 2>&1; tail -2 $O/r3r_tests.log
timeout 600 python tools/kernel_bench.py --only gather --local-corr-path 2 --scales s8,s4,s8u,s4u > $O/r3r_gather_p2.jsonl 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/r3r_gather_p2.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in d if k in ('visit','scale','flow','regime','ms','frac_hbm','C','r','side')})
PY
