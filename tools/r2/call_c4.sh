#!/bin/bash
# config 4 (RNA + ATAC, 2 M cells x 5 k genes + 50 k peaks) on one GPU; falls back to 500 k cells if the full size fails
mkdir -p gpurun_out
free -g | head -2
timeout 540 python bench.py --workload c4 --steps 30 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.log
rc=$?; echo "rc=$rc"; tail -4 gpurun_out/r2_bench_c4_n1.log | cut -c1-300
if [ $rc -ne 0 ]; then
  timeout 400 python bench.py --workload c4 --cells 500000 --steps 30 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_c4_n1_500k.json 2> gpurun_out/r2_bench_c4_n1_500k.log
  echo "rc=$?"; tail -4 gpurun_out/r2_bench_c4_n1_500k.log | cut -c1-300
  python tools/print_bench_line.py gpurun_out/r2_bench_c4_n1_500k.json 2>/dev/null | head -12
else
  python tools/print_bench_line.py gpurun_out/r2_bench_c4_n1.json 2>/dev/null | head -12
fi
nvidia-smi --query-gpu=memory.used --format=csv,noheader
