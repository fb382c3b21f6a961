#!/bin/bash
# ring GATv2 kernels, third version: parity tests of the kernels, then the microbenchmark
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gatv2_gpu.py -m gpu -x -q -s 2>&1 | grep -E "gatv2|passed|failed|Error|error" | tail -20 | tee gpurun_out/r2h_gatv2_tests.txt
timeout 200 python tools/bench_gatv2.py --iters 20 2>&1 | tail -4 | tee gpurun_out/r2h_gatv2_bench.txt
