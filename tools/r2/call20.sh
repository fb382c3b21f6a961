#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r2c20_gpu_suite.txt
echo "== bench graphs on"
timeout 900 python bench.py --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r2c20_bench.json 2> gpurun_out/r2c20_bench.log
echo "rc=$?"; tail -3 gpurun_out/r2c20_bench.log; python tools/print_bench_line.py gpurun_out/r2c20_bench.json 2>/dev/null | head -30
