import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_gpu_engine import _inputs
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer
torch.manual_seed(0)
B, H, W = 2, 160, 224
model = RetinaNet(80, "resnet50")
mk = lambda g: NativeTrainer(model=copy.deepcopy(model), device="cuda", batch=B, image_hw=(H, W), lr=1e-4, clipnorm=0.001, use_graph=g)
a, b, c = mk(True), mk(False), mk(False)
for step in range(7):
    img, gt, cnt, hw = _inputs(B, H, W, seed=step)
    if step % 2:
        gt, cnt = gt[:, :3], torch.full_like(cnt, 3)
    la = a.train_step(img, gt, cnt, hw).cpu(); lb = b.train_step(img, gt, cnt, hw).cpu(); lc = c.train_step(img, gt, cnt, hw).cpu()
    print(step, "graph", la.tolist(), "eager", lb.tolist(), "eager2", lc.tolist(),
          "dP graph-eager", (a.P - b.P).abs().max().item(), "dP eager-eager", (c.P - b.P).abs().max().item())
