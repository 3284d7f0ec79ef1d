#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 3 --master-addr 127.0.0.1 --master-port 29581 \
  -m moose_b200.party --mlp 65536 --fused --peer-memory --lanes ${1:-1} --ring-mib 512 --repeat 3 --trace 2>&1 | grep "^{" > gpurun_out/party_trace_l${1:-1}.json
python - <<PY
import json
d=json.loads(open('gpurun_out/party_trace_l${1:-1}.json').read().strip().splitlines()[-1])
print(d['seconds_per_run'], d['lanes'])
for i,p in enumerate(d['per_party']):
    for l,t in enumerate(p.get('trace_ms',[])):
        tot=sum(t.values()); wait=t.get('peer_wait',0)
        print('party',i,'lane',l,'total',round(tot,2),'wait',round(wait,2),'busy',round(tot-wait,2))
        print('   ',dict(list(t.items())[:14]))
PY
