#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_dp_exchange_gpu.py -x -q 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 tools/bench_dp_exchange.py 2>/dev/null | tee gpurun_out/r2n2c_dp_exchange.json
echo "== bench N=2 peer, graphs"
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 2 --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r2n2c_bench_peer.json 2> gpurun_out/r2n2c_bench_peer.log
echo "rc=$?"; python tools/print_bench_line.py gpurun_out/r2n2c_bench_peer.json 2>/dev/null | head -7
