#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/bench_inference_atlas.py --cells 300000 --check 2> gpurun_out/r2n2b_inf1.log | tee gpurun_out/r2n2b_inf_n1.json; tail -3 gpurun_out/r2n2b_inf1.log
for h in peer nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/bench_inference_atlas.py --cells 400000 --check --halo $h 2> gpurun_out/r2n2b_inf2_$h.log | tee gpurun_out/r2n2b_inf_n2_$h.json; tail -3 gpurun_out/r2n2b_inf2_$h.log
done
echo "== bench N=2 peer, graphs"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r2n2b_bench_peer.json 2> gpurun_out/r2n2b_bench_peer.log
echo "rc=$?"; tail -2 gpurun_out/r2n2b_bench_peer.log; python tools/print_bench_line.py gpurun_out/r2n2b_bench_peer.json 2>/dev/null | head -7
