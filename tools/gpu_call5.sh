#!/bin/bash
# profiles of one C2-shaped step with the CTA-pair GEMM: launch list (+DRAM bytes) and full-set capture
mkdir -p gpurun_out
bash tools/gpu_launchlist.sh
CMD="python tools/profile_step.py 100000 1"
KS='k_gemm_bf16|k_gatv2|k_nb_fused|k_adam|k_cast_bf16|k_col_sum|k_spmm|k_csr_gather|k_cosine|k_relu|k_masked_sddmm'
timeout 1200 ncu --set full --clock-control none --profile-from-start off -k regex:"$KS" \
   -o /tmp/r1f_step_full -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i /tmp/r1f_step_full.ncu-rep --page raw --csv > gpurun_out/r1f_step_full_raw.csv 2>/dev/null
# source-level capture of the two largest pair-GEMM launches
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'k_gemm_bf16' -s 12 -c 2 \
   -o gpurun_out/r1f_src_gemm_pair -f $CMD > gpurun_out/ncu_src_gemm.log 2>&1; echo "ncu src rc=$?"
ls -la gpurun_out | tail -8
