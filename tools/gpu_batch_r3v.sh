#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py tests/test_refparity_e2e_gpu.py -q -k "layernorm or vit or e2e or golden or config1 or patch" > $O/r3v_tests.log 2>&1; tail -3 $O/r3v_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $O/r3v_bench.log 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/r3v_bench.log'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['clocks']['sm_mhz'])
        print([(k['kernel'],k['dims'],k['avg_ms'],k['launches_per_step']) for k in d['kernels'] if 'layernorm' in k['kernel']])
PY
