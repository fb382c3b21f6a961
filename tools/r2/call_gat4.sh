#!/bin/bash
# ring GATv2 kernels: parity tests and microbenchmark under tight time limits (a hung kernel must not eat the budget)
mkdir -p gpurun_out
timeout -s KILL 50 python -m pytest tests/test_gatv2_gpu.py -m gpu -x -q -s 2>&1 | grep -E "gatv2|passed|failed|Error|error" | tail -20 | tee gpurun_out/r2m_gatv2_tests.txt
if grep -q "9 passed" gpurun_out/r2m_gatv2_tests.txt; then
  timeout -s KILL 40 python tools/bench_gatv2.py --iters 20 2>&1 | tail -4 | tee gpurun_out/r2m_gatv2_bench.txt
fi
