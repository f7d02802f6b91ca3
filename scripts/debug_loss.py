import sys, os, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.engine.library import LibraryTrainer
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer

torch.manual_seed(0)
B, H, W = 2, 800, 1333
g = torch.Generator().manual_seed(1)
img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g)
G = 12
xy = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.7, H * 0.7])
wh = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.3, H * 0.3]) + 24
gt = torch.cat([xy, xy + wh, torch.randint(0, 80, (B, G, 1), generator=g).float()], dim=2)
cnt = torch.full((B,), G, dtype=torch.int32)
hw = torch.tensor([[H, W]] * B, dtype=torch.int32)
model = RetinaNet(80, "resnet50")
m2 = copy.deepcopy(model)
nt = NativeTrainer(model=model, device="cuda", batch=B, image_hw=(H, W))
lib = LibraryTrainer(model=m2, device="cuda")
lib32 = LibraryTrainer(model=copy.deepcopy(m2), device="cuda", dtype=torch.float32)
ln, plan = nt.forward_loss(img, gt, cnt, hw)
with torch.no_grad():
    l16 = lib.loss(img.cuda(), gt.cuda(), cnt.cuda(), hw.cuda())
    l32 = lib32.loss(img.cuda(), gt.cuda(), cnt.cuda(), hw.cuda())
print("native  [tot, reg, cls]", ln.tolist())
print("lib bf16 [tot, cls, reg]", [float(v) for v in l16])
print("lib fp32 [tot, cls, reg]", [float(v) for v in l32])
for si in range(4):
    o = plan["acts"][si][-1]["o"]
    print("stage", si + 2, "native act absmax", o.float().abs().max().item(), "mean", o.float().abs().mean().item())
print("P3 absmax", plan["fpn"]["P"][0].float().abs().max().item(), "cls_o absmax", plan["cls_o"][0].float().abs().max().item())
for i in range(6):
    a = nt.train_step(img, gt, cnt, hw).tolist()
    b = lib.train_step(img, gt, cnt, hw).tolist()
    print(i, "native", [round(v, 3) for v in a], " lib", [round(v, 3) for v in b], "clip scale", nt.scal[1].item(), "gnorm", nt.scal[0].sqrt().item())
