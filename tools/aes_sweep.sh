#!/bin/bash
# PRF ceiling sweep (run on a B200): AES-128-CTR blocks/s of csrc/aes.cuh with NB independent blocks in flight per thread
# and TPB threads per CTA (one CTA per SM: the table takes the shared-memory window).  Binaries are built beforehand:
#   for tpb in 256 512 1024; do for nb in 1 2 3 4 6 8; do nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 \
#     -DAES_ABS_TABLE=1 -DTPB=$tpb -DNB=$nb -o tools/_aesb/b_${tpb}_${nb} tools/aes_bench.cu moose_b200/csrc/hostcrypto.cpp; done; done
mkdir -p gpurun_out
: > gpurun_out/aes_sweep.log
for b in $(ls tools/_aesb | sort -t_ -k2n -k3n); do tools/_aesb/$b >> gpurun_out/aes_sweep.log 2>&1; done
cat gpurun_out/aes_sweep.log
