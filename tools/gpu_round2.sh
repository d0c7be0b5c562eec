#!/bin/bash
# One GPU-box call for the round-2 evidence: parity tests (-s log), bench (both arms), CUPTI breakdown, GEMM trace, ncu launch list of one
# C2 step and one `--set full` capture of the top kernels (each only after the same command exited 0 without ncu).
TAG=${1:-r2u}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu_s_${TAG}.log 2>&1; echo "pytest exit $?"; tail -n 2 gpurun_out/pytest_gpu_s_${TAG}.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_${TAG}.log 2>&1; echo "bench exit $?"; tail -n 1 gpurun_out/bench_${TAG}.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_${TAG}.log 2>&1; echo "ref exit $?"; tail -n 1 gpurun_out/bench_ref_${TAG}.log | cut -c1-300
timeout 300 python tools/step_breakdown.py 16 train > gpurun_out/breakdown_${TAG}.log 2>&1; head -n 12 gpurun_out/breakdown_${TAG}.log | cut -c1-160
timeout 300 python tools/gemm_trace.py > gpurun_out/gemm_trace_${TAG}.log 2>&1; head -n 1 gpurun_out/gemm_trace_${TAG}.log
timeout 300 python tools/one_step.py 16 train 1 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_train_${TAG}.csv python tools/one_step.py 16 train 1 > gpurun_out/ncu_launch.log 2>&1
timeout 300 python tools/one_step.py 16 train 1 > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm_pair|pna_aggregate|pna_dpre|act_bwd' -s 120 -c 14 -o gpurun_out/prof_${TAG} python tools/one_step.py 16 train 1 > gpurun_out/ncu_full.log 2>&1
tail -n 2 gpurun_out/ncu_full.log; ls -la gpurun_out | tail -n 12
