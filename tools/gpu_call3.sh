#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/bench_gemm.py > gpurun_out/gemm_sweep.txt 2>&1; echo "gemm rc=$?"; cat gpurun_out/gemm_sweep.txt
CMD="python tools/profile_step.py 100000 1"
KS='k_gemm_bf16|k_gatv2|k_nb_fused|k_adam|k_cast_bf16|k_col_sum|k_spmm|k_csr_gather|k_cosine|k_relu'
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain2.log; exit 1; }
# (A) every hot kernel of one step, full metric set, no source: exported to csv on the box
timeout 1200 ncu --set full --clock-control none --profile-from-start off -k regex:"$KS" \
   -o /tmp/r1_step_full -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu A rc=$?"
ncu -i /tmp/r1_step_full.ncu-rep --page raw --csv > gpurun_out/r1_step_full_raw.csv 2>/dev/null
ls -la /tmp/r1_step_full.ncu-rep gpurun_out/r1_step_full_raw.csv
# (B) source-level captures of the top kernels
cap() { # name regex skip count
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"$2" -s $3 -c $4 \
     -o gpurun_out/r1_src_$1 -f $CMD > gpurun_out/ncu_src_$1.log 2>&1; echo "ncu B $1 rc=$?"; }
cap gemm 'k_gemm_bf16' 8 3
cap gatv2_bwd_dst 'k_gatv2_bwd_dst' 0 1
cap gatv2_fwd 'k_gatv2_fwd' 1 1
cap nb_fused 'k_nb_fused' 0 1
ls -la gpurun_out/*.ncu-rep; du -sh gpurun_out
