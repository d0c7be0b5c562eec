#!/bin/bash
# One GPU-box call: parity tests, bench, CUPTI breakdown, ncu launch list + full capture of the top kernels.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench.log
python tools/step_breakdown.py 16 train > gpurun_out/breakdown_train.log 2>&1; head -30 gpurun_out/breakdown_train.log
python tools/step_breakdown.py 16 test > gpurun_out/breakdown_test.log 2>&1; head -12 gpurun_out/breakdown_test.log
python tools/one_step.py 16 train 1 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_train.csv python tools/one_step.py 16 train 1 > gpurun_out/ncu_launch.log 2>&1
python tools/one_step.py 16 train 1 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gemm_tf32x3|pna_aggregate|pna_dpre|act_bwd|gp_ift|gp_fit' -s 60 -c 30 -o gpurun_out/prof_r1e python tools/one_step.py 16 train 1 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
