#!/bin/bash
# round 2, call 1: ring GATv2 kernels -- correctness (memcheck on the kernel test), full GPU suite, short bench
mkdir -p gpurun_out
echo "== gatv2 kernel test"
timeout 600 python -m pytest tests/test_gatv2_gpu.py -x -q 2>&1 | tail -15 | tee gpurun_out/r2c1_gatv2_test.txt
echo "== memcheck (gatv2 test)"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 1 python -m pytest tests/test_gatv2_gpu.py -x -q 2>&1 | tail -30 > gpurun_out/r2c1_memcheck.txt
echo "memcheck rc=$?"; tail -5 gpurun_out/r2c1_memcheck.txt
echo "== full gpu suite"
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r2c1_gpu_suite.txt
echo "== bench (short, no cpu baseline)"
timeout 900 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2c1_bench.json 2> gpurun_out/r2c1_bench.log
echo "bench rc=$?"; tail -3 gpurun_out/r2c1_bench.log
python tools/print_bench_line.py gpurun_out/r2c1_bench.json 2>/dev/null | head -40
