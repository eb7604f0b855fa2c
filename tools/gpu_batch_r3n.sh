#!/bin/bash
O=gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_fullsize_gpu.py -q -x -k "dwconv or fused" > $O/r3n_tests.log 2>&1; tail -3 $O/r3n_tests.log
timeout 300 python tools/kernel_bench.py --only dwconv --iters 20 > $O/r3n_dwconv.jsonl 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/r3n_dwconv.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print(d['visit'], d['C'], d['side'], d['ms'], d['frac_hbm'])
PY
