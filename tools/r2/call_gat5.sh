#!/bin/bash
# ring GATv2 kernels as shipped: ncu full capture of one forward + backward (tools/bench_gatv2.py --once)
mkdir -p gpurun_out
timeout -s KILL 70 ncu --set full --clock-control none --import-source on -k regex:k_gat -o gpurun_out/r2n_ncu_gatv2 -f python tools/bench_gatv2.py --once > gpurun_out/r2n_ncu_gatv2.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/r2n_ncu_gatv2.log
