"""Three arms per trial: graph (a), eager (b), eager (c), same init and inputs.  Is the graph-vs-eager drift larger
than the eager-vs-eager drift?  Prints per-step relative loss differences and final max |dP| for both pairs."""
import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_gpu_engine import _inputs
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer
B, H, W = 2, 160, 224
rel = lambda x, y: "%.1e" % ((x - y).abs() / y.abs()).max().item()
for trial in range(int(os.environ.get("TRIALS", 8))):
    torch.manual_seed(0)
    model = RetinaNet(80, "resnet50")
    mk = lambda g: NativeTrainer(model=copy.deepcopy(model), device="cuda", batch=B, image_hw=(H, W), lr=float(os.environ.get("LR", 1e-4)), clipnorm=float(os.environ.get("CLIP", 0.001)), use_graph=g)
    a, b, c = mk(True), mk(False), mk(False)
    ge, ee = [], []
    for step in range(6):
        batch = _inputs(B, H, W, seed=step)
        la, lb, lc = (t.train_step(*batch).cpu() for t in (a, b, c))
        ge.append(rel(la, lb)); ee.append(rel(lc, lb))
    torch.cuda.synchronize()
    print("trial", trial, "graph-eager", ge, "dP %.1e" % (a.P - b.P).abs().max().item(), "| eager-eager", ee,
          "dP %.1e" % (c.P - b.P).abs().max().item(), flush=True)
