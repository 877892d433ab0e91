#!/bin/bash
# round 2 evidence in one call: GPU suite, smoke, both bench arms, long-utterance stress, CLI end to end, ncu launch list,
# ncu --set full of the attention kernel (162- and 750-token shapes) and of the layer GEMMs at the bench batch size,
# compute-sanitizer logs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2e_pytest.log; tail -3 gpurun_out/r2e_pytest.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2e_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r2e_smoke.log
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2e_bench_reference.json 2> gpurun_out/r2e_bench_reference.err; echo "ref exit $?"; cut -c1-400 gpurun_out/r2e_bench_reference.json
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; echo "bench exit $?"; tail -2 gpurun_out/r2e_bench.err; cut -c1-700 gpurun_out/r2e_bench.json
timeout 600 python bench.py --long-stress --steps 3 > gpurun_out/r2e_long.json 2> gpurun_out/r2e_long.err; echo "long exit $?"; cut -c1-300 gpurun_out/r2e_long.json
timeout 900 python bench.py --cli-e2e 100000 > gpurun_out/r2e_cli_e2e.json 2> gpurun_out/r2e_cli_e2e.err; echo "cli exit $?"; tail -2 gpurun_out/r2e_cli_e2e.err | cut -c1-300; cut -c1-600 gpurun_out/r2e_cli_e2e.json
timeout 300 python tools/twin_latency.py 5 > gpurun_out/r2e_twin_latency.json 2>/dev/null; cut -c1-500 gpurun_out/r2e_twin_latency.json
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/r2e_kprof_bf16x3.txt 2>&1; head -16 gpurun_out/r2e_kprof_bf16x3.txt | cut -c1-140
# ---- ncu: launch list of the bench command (same flags minus the CPU legs), then full captures
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2e_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-bf16-mode --no-frontend > gpurun_out/r2e_ncu_launch.log 2>&1; echo "ncu launch list exit $?"; wc -l gpurun_out/r2e_launches.csv
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:attention_pp" -s 2 -c 1 -o gpurun_out/r2e_attn_pp_162 python tools/one_step.py 250 bf16x3 1 > gpurun_out/r2e_ncu_a.log 2>&1; tail -1 gpurun_out/r2e_ncu_a.log
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:attention_pp" -s 2 -c 1 -o gpurun_out/r2e_attn_pp_750 python tools/pp_one_long.py > gpurun_out/r2e_ncu_b.log 2>&1; tail -1 gpurun_out/r2e_ncu_b.log
timeout 600 ncu --set full --clock-control none -k "regex:gemm_kernel<256" -s 2 -c 4 -o gpurun_out/r2e_gemm_pair_bench python tools/one_step.py 370 bf16x3 1 > gpurun_out/r2e_ncu_c.log 2>&1; tail -1 gpurun_out/r2e_ncu_c.log
# ---- sanitizers
timeout 900 compute-sanitizer --tool memcheck python tools/sanitize_step.py bf16x3 > gpurun_out/r2e_memcheck_bf16x3.log 2>&1; tail -3 gpurun_out/r2e_memcheck_bf16x3.log | cut -c1-200
timeout 900 compute-sanitizer --tool memcheck python tools/sanitize_step.py bf16 > gpurun_out/r2e_memcheck_bf16.log 2>&1; tail -3 gpurun_out/r2e_memcheck_bf16.log | cut -c1-200
timeout 900 compute-sanitizer --tool racecheck python tools/sanitize_step.py bf16x3 > gpurun_out/r2e_racecheck.log 2>&1; tail -3 gpurun_out/r2e_racecheck.log | cut -c1-200
timeout 600 compute-sanitizer --tool synccheck python tools/sanitize_step.py bf16x3 > gpurun_out/r2e_synccheck.log 2>&1; tail -3 gpurun_out/r2e_synccheck.log | cut -c1-200
