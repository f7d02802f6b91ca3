#!/bin/bash
# A/B on one 8-GPU box: gradient scatter during backward (B200_AR_OVERLAP) on / off, per-GPU batch 16
cd "$(dirname "$0")/.."
run8() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 2000)) bench.py --gpus 8 --batch 16 --steps 60 --warmup 10 --no-e2e 2>/dev/null | grep '^{' | python -c "import json,sys; r=json.loads(sys.stdin.read()); print(round(r['ms_per_step'],3), 'ms', round(r['value'],1))"; }
for OV in 1 0; do echo -n "8 GPUs overlap=$OV: "; B200_AR_OVERLAP=$OV run8; done
