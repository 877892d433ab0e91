#!/bin/bash
mkdir -p gpurun_out
for bs in 1200 2400 3600 7200; do
  timeout 600 python bench.py --steps 4 --warmup 3 --batch-seconds $bs --no-cpu-baseline --no-bf16-mode > gpurun_out/bench_bs$bs.json 2> gpurun_out/bench_bs$bs.err; echo "bs $bs exit $?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_bs$bs.json'))
print($bs, round(d['ms_per_step'],2), round(d['value'],3), round(d['e2e']['value'],3), d['clocks']['sm_mhz'], d['gpu_launches'])
PY
done
