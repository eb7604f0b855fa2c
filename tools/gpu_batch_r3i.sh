#!/bin/bash
O=gpurun_out
timeout 600 python -m pytest tests/test_topk_gpu.py -q -x > $O/r3i_topk.log 2>&1; tail -15 $O/r3i_topk.log
timeout 200 python tools/topk_bench.py > $O/r3i_topk_bench.jsonl 2>&1; cat $O/r3i_topk_bench.jsonl
timeout 900 python -m pytest tests/test_kde_exact_gpu.py tests/test_e2e_gpu.py tests/test_refparity_e2e_gpu.py -q -x -k "sample or tiny or patch" > $O/r3i_tests.log 2>&1; tail -3 $O/r3i_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $O/r3i_bench.log 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/r3i_bench.log'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['clocks']['sm_mhz'])
PY
