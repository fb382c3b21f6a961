#!/bin/bash
# round 2, 8-GPU call: C2 bench (captured iterations + fused NVLink exchange), config-5 inference at 8.5 M cells,
# config-3 (8 samples x 500 k cells, sample covariate) bench
mkdir -p gpurun_out
free -g | head -2
echo "== C2 N=8"
timeout 360 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r2n8_bench_c2.json 2> gpurun_out/r2n8_bench_c2.log
echo "rc=$?"; python tools/print_bench_line.py gpurun_out/r2n8_bench_c2.json 2>/dev/null | head -8
echo "== C5 inference 8.5M"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 tools/bench_inference_atlas.py --cells 8500000 --check 2> gpurun_out/r2n8_inf.log | tee gpurun_out/r2n8_inference_c5.json
echo "rc=$?"; tail -2 gpurun_out/r2n8_inf.log | cut -c1-200
rm -f /dev/shm/nc_atlas_* /dev/shm/nc_bench_*
echo "== C3 N=8"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --workload c3 --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/r2n8_bench_c3.json 2> gpurun_out/r2n8_bench_c3.log
echo "rc=$?"; tail -2 gpurun_out/r2n8_bench_c3.log | cut -c1-200; python tools/print_bench_line.py gpurun_out/r2n8_bench_c3.json 2>/dev/null | head -8
rm -f /dev/shm/nc_atlas_* /dev/shm/nc_bench_*
free -g | head -2
