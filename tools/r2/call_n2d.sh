#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_dp_exchange_gpu.py tests/test_trainer_gpu.py -x -q -k "peer or cuda_graph" 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 tools/bench_dp_exchange.py --ctas 0,296 2>/dev/null | tee gpurun_out/r2n2d_dp_exchange.json
