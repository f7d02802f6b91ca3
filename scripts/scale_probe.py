"""Under torchrun: device-timed ms/step of the native engine with the overlapped and the plain allreduce (same process,
same box), plus one traced step (phases, scatter pieces).  Prints JSON lines on rank 0."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
H, W, G = 800, 1333, 12
g = torch.Generator().manual_seed(1 + rank)
sets = []
for _ in range(3):
    img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g)
    xy = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.7, H * 0.7])
    wh = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.3, H * 0.3]) + 24.0
    gt = torch.cat([xy, xy + wh, torch.randint(0, 80, (B, G, 1), generator=g).float()], dim=2)
    sets.append(tuple(t.to(dev) for t in (img, gt, torch.full((B,), G, dtype=torch.int32), torch.tensor([[H, W]] * B, dtype=torch.int32))))
nt = NativeTrainer("resnet50", 80, device=dev, batch=B, image_hw=(H, W), distributed=True)


def timed(n):
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(n):
        nt.train_step(*sets[i % 3])
    e.record(); torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e) / n], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item()


out = {"world": world, "batch": B}
for mode in ("1", "0", "1", "0"):
    os.environ["B200_AR_OVERLAP"] = mode
    nt._plans.clear()            # the captured graphs hold one of the two paths
    for i in range(6):
        nt.train_step(*sets[i % 3])
    out.setdefault("overlap_ms" if mode == "1" else "plain_ms", []).append(round(timed(steps), 4))
os.environ["B200_AR_OVERLAP"] = "1"
ph = nt.trace_step(*sets[0])
ph = {k: (round(v, 4) if isinstance(v, float) else v) for k, v in ph.items()}
os.environ["B200_AR_OVERLAP"] = "0"
ph0 = {k: round(v, 4) for k, v in nt.profile_step(*sets[0]).items()}
allph = [None] * world
dist.all_gather_object(allph, ph.get("allreduce"))
if rank == 0:
    out["trace_overlap"] = ph
    out["trace_plain"] = ph0
    out["allreduce_phase_ms_per_rank"] = allph
    print(json.dumps(out))
nt.comm.close()
dist.destroy_process_group()
