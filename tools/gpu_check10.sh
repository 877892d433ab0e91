#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_invariance.py tests/test_gpu_frontend.py -q > gpurun_out/pytest_inv.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_inv.log
tail -6 gpurun_out/pytest_inv.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print({k:d[k] for k in ('value','ms_per_step','e2e','clocks','gpu_launches')}); print(d.get('bf16_mode')); print(d['roofline'].get('executed_frac'))
PY
