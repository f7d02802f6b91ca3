"""Phase-level device timing + CPU launch time of one native training step (B=8, 800x1333)."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer

B, H, W = int(os.environ.get("B", 8)), 800, 1333
g = torch.Generator().manual_seed(1)
img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g).cuda()
G = 12
xy = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.7, H * 0.7])
wh = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.3, H * 0.3]) + 24
gt = torch.cat([xy, xy + wh, torch.randint(0, 80, (B, G, 1), generator=g).float()], dim=2).cuda()
cnt = torch.full((B,), G, dtype=torch.int32).cuda()
hw = torch.tensor([[H, W]] * B, dtype=torch.int32).cuda()
nt = NativeTrainer("resnet50", 80, device="cuda", batch=B, image_hw=(H, W))
steps = int(os.environ.get("STEPS", 3))
for _ in range(3):
    nt.train_step(img, gt, cnt, hw)
torch.cuda.synchronize()
if os.environ.get("PHASES", "1") == "1":
    t0 = time.perf_counter()
    for _ in range(steps):
        nt.train_step(img, gt, cnt, hw)
    t_cpu = (time.perf_counter() - t0) / steps
    torch.cuda.synchronize()
    t_all = (time.perf_counter() - t0) / steps
    print(json.dumps({"cpu_launch_ms_per_step": t_cpu * 1e3, "wall_ms_per_step": t_all * 1e3}))
    ph = nt.profile_step(img, gt, cnt, hw)
    print(json.dumps({k: round(v, 3) for k, v in ph.items()}))
    print("sum", sum(ph.values()))
else:
    for _ in range(steps):
        nt.train_step(img, gt, cnt, hw)
    torch.cuda.synchronize()
