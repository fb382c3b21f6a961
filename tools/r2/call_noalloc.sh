#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
echo "== alone"
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2g_noalloc_alone.txt
echo "== full suite"
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r2g_gpu_suite.txt
