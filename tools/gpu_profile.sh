#!/bin/bash
# ncu evidence for profiles/: launch list of one C2 training step + one `--set full` capture of the top kernels (kept small:
# gpurun merges at most 64 MiB back)
mkdir -p gpurun_out
python tools/one_step.py 16 train 1 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_train.csv python tools/one_step.py 16 train 1 > gpurun_out/ncu_launch.log 2>&1
python tools/one_step.py 16 train 1 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gemm_tf32x3|pna_aggregate|pna_dpre|act_bwd' -s 70 -c 14 -o gpurun_out/prof_r1e python tools/one_step.py 16 train 1 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log; ls -la gpurun_out
