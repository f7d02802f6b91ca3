#!/bin/bash
# Weak-scaling points for one GPU count: bench.py at per-GPU batch 1, 8 and 16 on N GPUs of this box (one process per
# GPU, torchrun, device-timed max over ranks), one JSON line each into gpurun_out/scaling_v5_N<N>.jsonl.  With N = 8 also
# the allreduce sweep.   usage: scripts/scaling_curve.sh N
set -u
N=${1:-1}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
O=gpurun_out/scaling_v5_N${N}.jsonl
: > $O
run() {  # extra bench args
  if [ "$N" = 1 ]; then timeout 400 python bench.py --gpus 1 "$@"
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 2000)) bench.py --gpus $N "$@"; fi
}
for B in 16 8 1; do
  S=100; [ $B = 1 ] && S=300; [ $B = 8 ] && S=150
  run --batch $B --steps $S --warmup 10 2> gpurun_out/scaling_N${N}_b${B}.err | grep '^{' >> $O
  tail -1 $O | cut -c1-200
done
if [ "$N" = 8 ]; then
  MAX_BYTES=$((256 << 20)) timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29431 scripts/bench_allreduce.py 2> gpurun_out/ar_sweep.err | grep '^{' > gpurun_out/allreduce_sweep_8gpu_v4.jsonl
  tail -4 gpurun_out/allreduce_sweep_8gpu_v4.jsonl | cut -c1-300
fi
