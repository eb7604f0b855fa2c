#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_fullsize_gpu.py tests/test_e2e_gpu.py -q -x -k "dwconv or fused or e2e or golden" > $O/r3l_tests.log 2>&1; tail -3 $O/r3l_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $O/r3l_bench.log 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/r3l_bench.log'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['frac'], d['roofline']['avg_launch_ms'])
        print([ (k['dims'].get('C'), k['dims'].get('h'), k['avg_ms']) for k in d['kernels'] if 'dwconv' in k['kernel']])
PY
