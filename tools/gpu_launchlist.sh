#!/bin/bash
# one profiled step: per-launch duration + DRAM bytes of every kernel
CMD="python tools/profile_step.py 100000 1"
timeout 300 $CMD > gpurun_out/plain_ll.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  --profile-from-start off --csv --log-file gpurun_out/step_launches.csv $CMD > gpurun_out/ncu_ll.log 2>&1
echo "rc=$?"; wc -l gpurun_out/step_launches.csv
