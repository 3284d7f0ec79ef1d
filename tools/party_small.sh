#!/bin/bash
# host-side cost of the party runtime: the same protocol on a batch so small that the GPUs have nothing to do
for b in 1024 8192; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 3 --master-addr 127.0.0.1 --master-port 29583 \
  -m moose_b200.party --mlp $b --fused --peer-memory --ring-mib 256 --repeat 10 2>&1 | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('batch', '$b', 'seconds_per_run', d['seconds_per_run'], d['per_party'][0])"
done
