#!/bin/bash
# round-1 re-entry call: GPU tests, smoke, bench (N=1), launch list with DRAM bytes, full-set capture of one step
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_n1.json | cut -c1-1500
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json | cut -c1-600
bash tools/gpu_launchlist.sh
CMD="python tools/profile_step.py 100000 1"
KS='k_gemm_bf16|k_gatv2|k_nb_fused|k_adam|k_cast_bf16|k_col_sum|k_spmm|k_csr_gather|k_cosine|k_relu|k_sddmm|k_attn'
timeout 1200 ncu --set full --clock-control none --profile-from-start off -k regex:"$KS" \
   -o /tmp/r1_step_full -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i /tmp/r1_step_full.ncu-rep --page raw --csv > gpurun_out/r1c_step_full_raw.csv 2>/dev/null
ls -la gpurun_out
