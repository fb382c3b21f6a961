#!/bin/bash
# round 2, call 2: ring GATv2 at every shape, gather tests without retry, full GPU suite, GATv2 microbench + ncu
mkdir -p gpurun_out
echo "== gatv2 + kernels tests"
timeout 600 python -m pytest tests/test_gatv2_gpu.py tests/test_kernels_gpu.py -x -q 2>&1 | tail -15 | tee gpurun_out/r2c2_kernel_tests.txt
echo "== gather tests x20"
for i in $(seq 1 20); do timeout 120 python -m pytest tests/test_kernels_gpu.py -x -q -k "gather or host_store" 2>&1 | tail -1; done | sort | uniq -c | tee gpurun_out/r2c2_gather_loop.txt
echo "== full gpu suite"
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r2c2_gpu_suite.txt
echo "== gatv2 microbench"
timeout 300 python tools/bench_gatv2.py 2>&1 | tee gpurun_out/r2c2_gatv2_bench.txt
timeout 300 python tools/bench_gatv2.py --generic 2>&1 | tee -a gpurun_out/r2c2_gatv2_bench.txt
timeout 300 python tools/bench_gatv2.py --widths 300,100,300,100 --n-src 6000 2>&1 | tee -a gpurun_out/r2c2_gatv2_bench.txt
echo "== ncu gatv2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gat -o gpurun_out/r2c2_gatv2 -f python tools/bench_gatv2.py --once > gpurun_out/r2c2_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r2c2_ncu.log
