#!/bin/bash
mkdir -p gpurun_out
echo "== exchange microbench N=8"
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 tools/bench_dp_exchange.py --ctas 0,74 2>/dev/null | tee gpurun_out/r2n8b_dp_exchange.json
echo "== C2 N=8 peer"
timeout 330 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 8 --steps 200 --warmup 20 --no-cpu-baseline --epoch > gpurun_out/r2n8b_bench_c2.json 2> gpurun_out/r2n8b_bench_c2.log
echo "rc=$?"; python tools/print_bench_line.py gpurun_out/r2n8b_bench_c2.json 2>/dev/null | head -8
python -c "
import json; d=json.load(open('gpurun_out/r2n8b_bench_c2.json')); print('epoch', d.get('epoch')); print('host', d.get('host_enqueue_ms_per_step'), d.get('cuda_graphs'))"
