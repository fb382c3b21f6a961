#!/bin/bash
# round 2, call 3: KATs on the kernels, C1-scale parity (fp32 / bf16 vs exact and bf16-rounding oracle), gather tests
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kat.py tests/test_kernels_gpu.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r2c3_kat.txt
timeout 1500 python -m pytest tests/test_c1_parity_gpu.py -x -q -s 2>&1 | tail -40 | tee gpurun_out/r2c3_c1_parity.txt
