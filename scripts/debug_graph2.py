import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_gpu_engine import _inputs
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer
B, H, W = 2, 160, 224
pin = os.environ.get("PIN", "1") == "1"
for trial in range(int(os.environ.get("TRIALS", 5))):
    torch.manual_seed(0)
    model = RetinaNet(80, "resnet50")
    mk = lambda g: NativeTrainer(model=copy.deepcopy(model), device="cuda", batch=B, image_hw=(H, W), lr=1e-4, clipnorm=0.001, use_graph=g)
    a, b = mk(True), mk(False)
    out = []
    for step in range(5):
        img, gt, cnt, hw = _inputs(B, H, W, seed=step)
        batch = [t.pin_memory() for t in (img, gt, cnt, hw)] if pin else [img, gt, cnt, hw]
        la = a.train_step(*batch).cpu(); lb = b.train_step(img, gt, cnt, hw).cpu()
        out.append("%.1e" % ((la - lb).abs() / lb.abs()).max().item())
    print("trial", trial, "rel diff per step", out, "dP %.1e" % (a.P - b.P).abs().max().item(), flush=True)
