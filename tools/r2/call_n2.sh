#!/bin/bash
# round 2, 2-GPU call: peer-memory exchange test, C2 bench at N = 2 with the fused exchange and with NCCL
mkdir -p gpurun_out
nvidia-smi topo -m 2>/dev/null | head -6
timeout 600 python -m pytest tests/test_dp_exchange_gpu.py -x -q 2>&1 | tail -15 | tee gpurun_out/r2n2_dp_test.txt
for mode in peer nccl; do
  echo "== bench N=2 exchange=$mode"
  NC_DP_EXCHANGE=$mode timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus 2 --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r2n2_bench_$mode.json 2> gpurun_out/r2n2_bench_$mode.log
  echo "rc=$?"; tail -2 gpurun_out/r2n2_bench_$mode.log
  python tools/print_bench_line.py gpurun_out/r2n2_bench_$mode.json 2>/dev/null | head -12
done
