#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r2c16_gpu_suite.txt
echo "== bench graphs on"
timeout 900 python bench.py --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r2c16_bench_graphs.json 2> gpurun_out/r2c16_bench_graphs.log
echo "rc=$?"; tail -3 gpurun_out/r2c16_bench_graphs.log; python tools/print_bench_line.py gpurun_out/r2c16_bench_graphs.json 2>/dev/null | head -8
echo "== bench graphs off"
NC_CUDA_GRAPHS=0 timeout 900 python bench.py --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r2c16_bench_eager.json 2> gpurun_out/r2c16_bench_eager.log
echo "rc=$?"; python tools/print_bench_line.py gpurun_out/r2c16_bench_eager.json 2>/dev/null | head -4
