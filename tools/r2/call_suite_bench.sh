#!/bin/bash
# full GPU suite + default-shape bench line without the CPU legs
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2k_gpu_suite.txt
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2k_bench_c2_n1.json 2> gpurun_out/r2k_bench_c2_n1.log
echo "rc=$?"; tail -2 gpurun_out/r2k_bench_c2_n1.log | cut -c1-300; python tools/print_bench_line.py gpurun_out/r2k_bench_c2_n1.json 2>/dev/null | head -30
