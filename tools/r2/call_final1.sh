#!/bin/bash
# round 2, final single-GPU evidence: full GPU suite, default bench line (with CPU baseline + parity + real epoch),
# reference arm, ncu launch list of one training iteration and full-metric captures of the top kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4 | tee gpurun_out/r2f_gpu_suite.txt
echo "== smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
echo "== default bench"
timeout 900 python bench.py --epoch > gpurun_out/r2f_bench_c2_n1.json 2> gpurun_out/r2f_bench_c2_n1.log
echo "rc=$?"; tail -2 gpurun_out/r2f_bench_c2_n1.log; python tools/print_bench_line.py gpurun_out/r2f_bench_c2_n1.json 2>/dev/null | head -14
echo "== reference arm"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 --ref-wall-s 60 > gpurun_out/r2f_bench_c2_reference_arm.json 2> gpurun_out/r2f_ref.log
echo "rc=$?"; cut -c1-400 gpurun_out/r2f_bench_c2_reference_arm.json
echo "== ncu launch list (one eager iteration)"
NC_CUDA_GRAPHS=0 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r2f_launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2f_ncu_list.log 2>&1
echo "rc=$?"; wc -l gpurun_out/r2f_launches_c2.csv
echo "== ncu full: gatv2 ring kernels"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_gat -o gpurun_out/r2f_ncu_gatv2 -f python tools/bench_gatv2.py --once > gpurun_out/r2f_ncu_gatv2.log 2>&1
echo "rc=$?"
echo "== ncu full: GEMM launches of one eager iteration"
NC_CUDA_GRAPHS=0 timeout 600 ncu --set full --clock-control none -k regex:k_gemm_bf16 --launch-skip 80 -c 13 -o gpurun_out/r2f_ncu_gemm -f python bench.py --steps 1 --warmup 8 --no-cpu-baseline --no-e2e > gpurun_out/r2f_ncu_gemm.log 2>&1
echo "rc=$?"
