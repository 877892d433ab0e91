#!/bin/bash
# round-2 closing evidence in one 1-GPU call: GPU suite, smoke, headline bench, CLI end to end, per-kernel breakdown,
# ncu launch list of the bench command, ncu --set full of the front-end kernels and the CAM gate kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2f_pytest.log; tail -3 gpurun_out/r2f_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/r2f_smoke.log | cut -c1-200
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench exit $?"; tail -2 gpurun_out/r2f_bench.err | cut -c1-200; grep "^{" gpurun_out/r2f_bench.json | cut -c1-600
timeout 600 python bench.py --cli-e2e 100000 > gpurun_out/r2f_cli_e2e.json 2> gpurun_out/r2f_cli_e2e.err; echo "cli exit $?"; tail -2 gpurun_out/r2f_cli_e2e.err | cut -c1-300; cut -c1-2600 gpurun_out/r2f_cli_e2e.json
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/r2f_kprof_bf16x3.txt 2>&1; grep -v Warn gpurun_out/r2f_kprof_bf16x3.txt | head -22 | cut -c1-140
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2f_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-bf16-mode --no-frontend > gpurun_out/r2f_ncu_launch.log 2>&1; echo "ncu launch list exit $?"; wc -l gpurun_out/r2f_launches.csv
timeout 400 ncu --set full --clock-control none --import-source on -k "regex:fbank2_kernel|logmel_kernel|cmn_sub_kernel|logmel_finish_kernel" -s 4 -c 4 -o gpurun_out/r2f_frontend python tools/fe_one.py 1024 2 > gpurun_out/r2f_ncu_fe.log 2>&1; tail -1 gpurun_out/r2f_ncu_fe.log | cut -c1-200
timeout 400 ncu --set full --clock-control none --import-source on -k "regex:stem_kernel|fsmn_kernel" -c 2 -o gpurun_out/r2f_small python tools/one_step.py 250 bf16x3 1 > gpurun_out/r2f_ncu_small.log 2>&1; tail -1 gpurun_out/r2f_ncu_small.log | cut -c1-200
timeout 400 ncu --set full --clock-control none --import-source on -k "regex:cam_ctx_gate_kernel" -s 10 -c 1 -o gpurun_out/r2f_gate python tools/one_step.py 250 bf16x3 1 > gpurun_out/r2f_ncu_gate.log 2>&1; tail -1 gpurun_out/r2f_ncu_gate.log | cut -c1-200
ls -la gpurun_out | head -30
