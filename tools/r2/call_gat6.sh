#!/bin/bash
# forward variant with two rows in flight per warp (NC_GAT_FWD_PIPE=1): parity tests + microbenchmark, tight limits
mkdir -p gpurun_out
export NC_GAT_FWD_PIPE=1
timeout -s KILL 22 python -m pytest tests/test_gatv2_gpu.py -m gpu -x -q -s 2>&1 | grep -E "gatv2|passed|failed|Error|error" | tail -12 | tee gpurun_out/r2o_gatv2_pipe_tests.txt
if grep -q "9 passed" gpurun_out/r2o_gatv2_pipe_tests.txt; then
  timeout -s KILL 12 python tools/bench_gatv2.py --iters 20 2>&1 | tail -2 | tee gpurun_out/r2o_gatv2_pipe_bench.txt
fi
