#!/bin/bash
# quick GPU check: gemm/graph/model parity tests, then a short bench and the CUPTI breakdown
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_graph.py tests/test_gpu_model.py tests/test_gpu_gp.py -m gpu -x -q 2>&1 | tail -15
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-meta-test > gpurun_out/bench_quick.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench_quick.log | cut -c1-400
timeout 600 python tools/step_breakdown.py 16 train > gpurun_out/breakdown_quick.log 2>&1; head -24 gpurun_out/breakdown_quick.log
