#!/bin/bash
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > $O/final4_tests.log 2>&1; tail -2 $O/final4_tests.log
timeout 300 python __graft_entry__.py smoke > $O/final4_smoke.log 2>&1; tail -2 $O/final4_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $O/final4_bench.log 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/final4_bench.log'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['frac'])
PY
