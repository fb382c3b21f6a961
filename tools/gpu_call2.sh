#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_gemm.py > gpurun_out/gemm_sweep.txt 2>&1; echo "gemm rc=$?"; cat gpurun_out/gemm_sweep.txt
CMD="python tools/profile_step.py 100000 1"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on --profile-from-start off \
   -k regex:'k_gemm_bf16|k_gatv2|k_nb_fused|k_adam|k_cast_bf16|k_col_sum|k_spmm|k_csr_gather|k_cosine|k_relu' \
   -o gpurun_out/r1_step_full -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/*.ncu-rep
