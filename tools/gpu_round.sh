#!/bin/bash
# One GPU-box call: parity tests, bench (both arms), CUPTI breakdown, ncu launch list + full capture of the top kernels.
TAG=${1:-r1g}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench.log
timeout 400 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?"; tail -1 gpurun_out/bench_ref.log
timeout 300 python tools/step_breakdown.py 16 train > gpurun_out/breakdown_train.log 2>&1; head -30 gpurun_out/breakdown_train.log
timeout 300 python tools/one_step.py 16 train 1 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_train_${TAG}.csv python tools/one_step.py 16 train 1 > gpurun_out/ncu_launch.log 2>&1
timeout 300 python tools/one_step.py 16 train 1 > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm_tf32x3|pna_aggregate|pna_dpre|act_bwd' -s 70 -c 16 -o gpurun_out/prof_${TAG} python tools/one_step.py 16 train 1 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
