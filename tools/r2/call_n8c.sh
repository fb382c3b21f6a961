#!/bin/bash
# round 2, last 8-GPU call: C2 with a real epoch (sharded validation, ranks pinned to disjoint cores), C3 with the
# captured iteration (covariate embeddings)
mkdir -p gpurun_out
echo "== C2 N=8"
timeout 330 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 200 --warmup 20 --no-cpu-baseline --epoch > gpurun_out/r2n8c_bench_c2.json 2> gpurun_out/r2n8c_bench_c2.log
echo "rc=$?"; python tools/print_bench_line.py gpurun_out/r2n8c_bench_c2.json 2>/dev/null | head -8
python -c "
import json; d=json.load(open('gpurun_out/r2n8c_bench_c2.json')); print('epoch', d.get('epoch')); print('host', d.get('host_enqueue_ms_per_step'), d.get('cuda_graphs'))"
rm -f /dev/shm/nc_bench_*
echo "== C3 N=8"
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --workload c3 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e > gpurun_out/r2n8c_bench_c3.json 2> gpurun_out/r2n8c_bench_c3.log
echo "rc=$?"; tail -2 gpurun_out/r2n8c_bench_c3.log | cut -c1-200; python tools/print_bench_line.py gpurun_out/r2n8c_bench_c3.json 2>/dev/null | head -8
python -c "
import json; d=json.load(open('gpurun_out/r2n8c_bench_c3.json')); print('host', d.get('host_enqueue_ms_per_step'), d.get('cuda_graphs'))"
rm -f /dev/shm/nc_bench_*
