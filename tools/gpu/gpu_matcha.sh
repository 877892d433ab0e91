#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_matcha.py tests/test_gpu_frontend.py -q > gpurun_out/pytest_matcha.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_matcha.log
tail -25 gpurun_out/pytest_matcha.log
timeout 300 python tools/bench_frontend.py > gpurun_out/frontend_sweep.jsonl 2> gpurun_out/frontend_sweep.err; tail -3 gpurun_out/frontend_sweep.err; python - <<'PY'
import json
for l in open('gpurun_out/frontend_sweep.jsonl'):
    d=json.loads(l); print(d['batch'], {k:(round(v['ms'],3), round(v['frac_of_measured_hbm'],4)) for k,v in d.items() if isinstance(v,dict)})
PY
