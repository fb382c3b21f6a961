#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gat_ring -o gpurun_out/r2c22_gatv2 -f python tools/bench_gatv2.py --once > gpurun_out/r2c22_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r2c22_ncu.log
