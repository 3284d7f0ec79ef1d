#!/bin/bash
# PRF ceiling sweep (run on a B200): AES-128-CTR blocks/s of csrc/aes.cuh with NB independent blocks in flight per thread
# and TPB threads per CTA (one CTA per SM: the table takes the shared-memory window).  Output: gpurun_out/aes_sweep.log
set -e
mkdir -p gpurun_out /tmp/aesb
: > gpurun_out/aes_sweep.log
for tpb in 256 512 1024; do
  for nb in 1 2 3 4 6 8; do
    if [ $((tpb * nb)) -gt 4096 ]; then continue; fi
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DAES_ABS_TABLE=1 -DTPB=$tpb -DNB=$nb -o /tmp/aesb/b_${tpb}_${nb} \
      tools/aes_bench.cu moose_b200/csrc/hostcrypto.cpp 2>/dev/null || { echo "tpb=$tpb nb=$nb: compile failed" >> gpurun_out/aes_sweep.log; continue; }
    /tmp/aesb/b_${tpb}_${nb} >> gpurun_out/aes_sweep.log 2>&1 || true
  done
done
cat gpurun_out/aes_sweep.log
