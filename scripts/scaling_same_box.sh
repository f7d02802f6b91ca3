#!/bin/bash
# Same-box weak-scaling ratio: N = 1 and N = 8 back to back on ONE 8-GPU box (box-to-box spread of the 1-GPU step is
# +-2 %, more than the effect being measured), per-GPU batch 16 and 1.  Lines go to gpurun_out/scaling_same_box_v5.jsonl.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
O=gpurun_out/scaling_same_box_v5.jsonl
: > $O
for B in 16 1; do
  S=60; [ $B = 1 ] && S=300
  timeout 300 python bench.py --gpus 1 --batch $B --steps $S --warmup 10 --no-e2e 2>/dev/null | grep '^{' >> $O
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 2000)) bench.py --gpus 8 --batch $B --steps $S --warmup 10 --no-e2e 2>/dev/null | grep '^{' >> $O
done
python - <<'PY'
import json
rows = [json.loads(l) for l in open("gpurun_out/scaling_same_box_v5.jsonl")]
for a, b in zip(rows[0::2], rows[1::2]):
    print(f"per-GPU batch {a['config']['per_gpu_batch']}: 1 GPU {a['value']:.1f} img/s ({a['ms_per_step']:.3f} ms), 8 GPUs {b['value']:.1f} ({b['ms_per_step']:.3f} ms), efficiency {b['value'] / 8 / a['value']:.3f}")
PY
