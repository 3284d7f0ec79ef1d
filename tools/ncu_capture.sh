#!/bin/bash
# Nsight Compute captures of one bench step per workload (run on a B200, after the plain run has exited 0):
#   launch list (durations) of every kernel of the step, and a --set full report of the step (LogReg: of the kernels
#   the MLP step does not have).  Outputs under gpurun_out/: <w>_launches.csv, <w>_full_raw.csv (the .ncu-rep is
#   exported to CSV on the box and dropped: it exceeds what gpurun brings back).
set -x
mkdir -p gpurun_out
for w in "$@"; do
  timeout 300 python tools/profile_step.py --workload $w > gpurun_out/${w}_plain.log 2>&1 || { echo "plain run of $w failed"; tail -5 gpurun_out/${w}_plain.log; continue; }
  timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${w}_launches.csv \
    python tools/profile_step.py --workload $w > gpurun_out/${w}_ncu_list.log 2>&1
  filter=""
  if [ "$w" = "logreg" ]; then filter="--kernel-name regex:inject_sum|pow2_bits|mul_trunc|prefix_or|poly_combine|rep_dot_skinny|planes_gather|replicate|adder_level|adder_init|rep_mul|trunc_pr|split_planes|ew_same3 -c 120"; fi
  timeout 1500 ncu --profile-from-start off --set full --clock-control none --import-source on $filter -f -o /tmp/${w}_full \
    python tools/profile_step.py --workload $w > gpurun_out/${w}_ncu_full.log 2>&1
  ncu -i /tmp/${w}_full.ncu-rep --page raw --csv > gpurun_out/${w}_full_raw.csv 2>/dev/null
  ls -la /tmp/${w}_full.ncu-rep gpurun_out/${w}_full_raw.csv
done
