#!/bin/bash
mkdir -p gpurun_out
DEV=$PWD/fangyan_tts_b200/libfangyan_b200_dev.so
[ -f $DEV ] && export FY_LIB=$DEV
timeout 900 python -m pytest tests/test_gpu_frontend.py tests/test_gpu_golden.py tests/test_gpu_invariance.py tests/test_gpu_campplus.py -x -q 2>&1 | tail -5 | cut -c1-300
timeout 300 python tools/bench_frontend.py > gpurun_out/r2_frontend_sweep_v2.jsonl 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/r2_frontend_sweep_v2.jsonl"):
    if not l.startswith("{"): continue
    d = json.loads(l)
    print(d["batch"], {k: (round(d[k]["ms"], 3), round(100 * d[k]["frac_of_measured_hbm"], 2)) for k in ("fbank_cmn", "logmel")})
PY
