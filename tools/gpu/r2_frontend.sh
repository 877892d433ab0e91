#!/bin/bash
# front-end second generation: parity / golden / invariance tests and the batch sweep, against FY_FRONTEND_V1=1
mkdir -p gpurun_out
DEV=$PWD/fangyan_tts_b200/libfangyan_b200_dev.so
[ -f $DEV ] && export FY_LIB=$DEV
echo "lib: ${FY_LIB:-default}"
timeout 900 python -m pytest tests/test_gpu_frontend.py tests/test_gpu_golden.py tests/test_gpu_invariance.py tests/test_gpu_campplus.py -x -q 2>&1 | tail -8 | cut -c1-300
timeout 300 python tools/bench_frontend.py > gpurun_out/r2_frontend_sweep_v2.jsonl 2>&1
FY_FRONTEND_V1=1 timeout 300 python tools/bench_frontend.py > gpurun_out/r2_frontend_sweep_v1.jsonl 2>&1
python - <<'PY'
import json
for tag in ("v1", "v2"):
    for l in open(f"gpurun_out/r2_frontend_sweep_{tag}.jsonl"):
        if not l.startswith("{"): continue
        d = json.loads(l)
        print(tag, d["batch"], {k: (round(d[k]["ms"], 3), round(100 * d[k]["frac_of_measured_hbm"], 2)) for k in ("fbank_cmn", "logmel")})
PY
