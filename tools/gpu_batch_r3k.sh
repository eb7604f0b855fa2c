#!/bin/bash
O=gpurun_out
timeout 600 python -m pytest tests/test_topk_gpu.py -q -x > $O/r3k_topk.log 2>&1; tail -8 $O/r3k_topk.log
timeout 200 python tools/topk_bench.py > $O/r3k_topk_bench.jsonl 2>&1; cat $O/r3k_topk_bench.jsonl
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"select_pass|count_kernel|write_kernel|all_keys|RadixSort" -c 40 --csv --log-file $O/r3k_topk_launches.csv python tools/topk_bench.py > $O/r3k_ncu.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > $O/r3k_bench.log 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/r3k_bench.log'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['clocks']['sm_mhz'])
        print([ (k['dims'], k['avg_ms']) for k in d['kernels'] if k['kernel']=='roma_topk_desc'])
PY
