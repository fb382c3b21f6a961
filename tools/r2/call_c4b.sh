#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_vgpgae_gpu.py tests/test_trainer_gpu.py tests/test_golden.py -q -x 2>&1 | tail -3
timeout 540 python bench.py --workload c4 --steps 30 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_c4_n1_b.json 2> gpurun_out/r2_bench_c4_n1_b.log
echo "rc=$?"; python tools/print_bench_line.py gpurun_out/r2_bench_c4_n1_b.json 2>/dev/null | head -10
