#!/bin/bash
# last N = 1 line of round 2 with the CPU legs (bounded), as the driver runs it
mkdir -p gpurun_out
timeout -s KILL 115 python bench.py --cpu-steps 3 > gpurun_out/r2n_bench_c2_n1.json 2> gpurun_out/r2n_bench_c2_n1.log
echo "rc=$?"; tail -2 gpurun_out/r2n_bench_c2_n1.log | cut -c1-300; python tools/print_bench_line.py gpurun_out/r2n_bench_c2_n1.json 2>/dev/null | head -12
