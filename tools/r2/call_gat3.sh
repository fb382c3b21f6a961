#!/bin/bash
# ring GATv2 kernels: parity tests, microbenchmark, per-launch durations (ncu, time only)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gatv2_gpu.py -m gpu -x -q -s 2>&1 | grep -E "gatv2|passed|failed|Error|error" | tail -20 | tee gpurun_out/r2l_gatv2_tests.txt
timeout 200 python tools/bench_gatv2.py --iters 20 2>&1 | tail -4 | tee gpurun_out/r2l_gatv2_bench.txt
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.per_cycle_active --clock-control none -k regex:k_gat --csv --log-file gpurun_out/r2l_gatv2_launches.csv python tools/bench_gatv2.py --once > /dev/null 2>&1
grep -v "^==" gpurun_out/r2l_gatv2_launches.csv | cut -d, -f5,13- | cut -c1-200 | tail -16
