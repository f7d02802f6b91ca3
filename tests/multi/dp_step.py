"""2-rank data-parallel native step == single-process step on the concatenated batch (same clip semantics
are NOT expected: clipping is per-rank, so compare with clipnorm disabled)."""
import os, sys
os.environ.setdefault("B200_AR_OVERLAP", "1")     # exercised here at 2 ranks; the default switches it on from 4 ranks up
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(1234 + rank)            # different init per rank: broadcast must fix it
model = RetinaNet(80, "resnet50")
B, H, W = 1, 160, 224
g = torch.Generator().manual_seed(7)
img = torch.randint(0, 256, (world * B, H, W, 3), dtype=torch.uint8, generator=g)
G = 4
xy = torch.rand(world * B, G, 2, generator=g) * torch.tensor([W * 0.6, H * 0.6])
wh = torch.rand(world * B, G, 2, generator=g) * 60 + 20
gt = torch.cat([xy, xy + wh, torch.randint(0, 80, (world * B, G, 1), generator=g).float()], dim=2)
cnt = torch.full((world * B,), G, dtype=torch.int32)
hw = torch.tensor([[H, W]] * (world * B), dtype=torch.int32)
nt = NativeTrainer(model=model, device=dev, batch=B, image_hw=(H, W), distributed=True, clipnorm=0.0, lr=1e-4)
p0 = nt.P.clone()
chk = p0.clone(); dist.broadcast(chk, 0)
assert torch.equal(chk, p0), "initial broadcast did not equalise the ranks"
sl = slice(rank * B, (rank + 1) * B)
loss = nt.train_step(img[sl], gt[sl], cnt[sl], hw[sl])
torch.cuda.synchronize()
gavg = nt.Gavg[:nt.n_train].clone()
ref = gavg.clone(); dist.broadcast(ref, 0)
assert torch.equal(ref, gavg), "averaged gradients differ between ranks"
# the overlapped (scatter during backward + staged reduce) result equals a plain allreduce of the same bucket
assert nt._overlap(), "B200_AR_OVERLAP=1 must select the overlapped allreduce"
plain = torch.zeros(nt.n_train + 64, device=dev)
nt.comm.allreduce_avg(nt.G, plain, nt.n_train, prescale=True, algo=2)
torch.cuda.synchronize()
d = (plain[:nt.n_train] - gavg).abs().max().item()
assert d <= 1e-6 * (1 + gavg.abs().max().item()), f"overlapped allreduce differs from the plain one by {d}"
p1 = nt.P.clone(); chk = p1.clone(); dist.broadcast(chk, 0)
assert torch.equal(chk, p1), "parameters diverged after one step"
assert not torch.equal(p1, p0)
for _ in range(3):
    loss = nt.train_step(img[sl], gt[sl], cnt[sl], hw[sl])
assert torch.isfinite(loss).all()
# the overlapped allreduce must be visible in the Horovod-style timeline: scatter pieces beside the backward pass
import json, tempfile
import batchai_retinanet_horovod_coco_b200.parallel as hvd
hvd.init()
tl = os.path.join(tempfile.gettempdir(), f"b200_tl_{os.environ.get('MASTER_PORT', '0')}.json")
if rank == 0:
    hvd.timeline_start(tl)
ph = nt.trace_step(img[sl], gt[sl], cnt[sl], hw[sl], tag="dp_step")
assert ph["allreduce"] > 0 and len(ph["grad_scatter_pieces"]) >= 4, ph
hvd.timeline_stop()
if rank == 0:
    names = {e["name"] for e in json.load(open(tl))}
    assert {"GPU_GRAD_SCATTER", "GPU_ALLREDUCE", "GPU_BACKWARD", "GPU_FORWARD"} <= names, names
    os.remove(tl)
nt.check_comm()
nt.comm.close()
hvd.shutdown()
if rank == 0:
    print("DP_OK", loss.tolist(), {k: round(v, 3) for k, v in ph.items() if isinstance(v, float)})
dist.destroy_process_group()
