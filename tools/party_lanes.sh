#!/bin/bash
# party-per-GPU MLP (config 5) over the peer-memory transport for several lane counts; one JSON line each
export TORCH_NCCL_SHOW_EAGER_INIT_P2P_SERIALIZATION_WARNING=false
B=${B:-65536}
mkdir -p gpurun_out
for L in "$@"; do
  timeout 250 python -m torch.distributed.run --nnodes=1 --nproc-per-node 3 --master-addr 127.0.0.1 --master-port $((29520 + L)) \
    -m moose_b200.party --mlp $B --fused --peer-memory --lanes $L --repeat 5 2>&1 | grep '^{"mode"' | tee -a gpurun_out/party_lanes.log |
    python -c 'import sys, json
for l in sys.stdin:
    d = json.loads(l); print({k: d[k] for k in ("lanes", "seconds_per_run", "rows_per_s", "max_abs_err_first_rows", "payload_bytes_per_run")})'
done
