#!/bin/bash
# A/B on one 2-GPU box: CTA pairs x allreduce overlap, per-GPU batch 16 (is the multi-GPU overhead the side-stream
# scatter competing with 2-CTA clusters for SMs?)
cd "$(dirname "$0")/.."
run2() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 2000)) bench.py --gpus 2 --batch 16 --steps 60 --warmup 10 --no-e2e 2>/dev/null | grep '^{' | python -c "import json,sys; r=json.loads(sys.stdin.read()); print(round(r['ms_per_step'],3), 'ms', round(r['value'],1))"; }
echo -n "1 GPU: "; timeout 300 python bench.py --gpus 1 --batch 16 --steps 60 --warmup 10 --no-e2e 2>/dev/null | grep '^{' | python -c "import json,sys; r=json.loads(sys.stdin.read()); print(round(r['ms_per_step'],3), 'ms', round(r['value'],1))"
for P in 1 0; do for OV in 1 0; do echo -n "2 GPUs pairs=$P overlap=$OV: "; B200_CONV_PAIRS=$P B200_AR_OVERLAP=$OV run2; done; done
