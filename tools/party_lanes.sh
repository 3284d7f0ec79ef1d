#!/bin/bash
# Config 5 on three GPUs (gpurun --gpus 4): party runtime over the peer-memory transport with 1 / 2 / 4 lanes.
mkdir -p gpurun_out
: > gpurun_out/party_lanes.log
for cfg in "1 0" "2 1" "2 0" "4 3" "3 2"; do
  set -- $cfg
  echo "# lanes=$1 reserve_sms=$2" >> gpurun_out/party_lanes.log
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 3 --master-addr 127.0.0.1 --master-port 2957$1 \
    -m moose_b200.party --mlp 65536 --fused --peer-memory --lanes $1 --reserve-sms $2 --ring-mib 512 --repeat 5 2>&1 | grep "^{" | \
    python -c "import sys,json; [print(json.dumps({k:v for k,v in json.loads(l).items() if k!='per_party'})) for l in sys.stdin]" >> gpurun_out/party_lanes.log
done
cat gpurun_out/party_lanes.log
