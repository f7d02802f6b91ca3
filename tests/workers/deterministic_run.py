"""B200_DETERMINISTIC=1: two trainers from the same state follow bit-identical trajectories, and CUDA-graph replay
follows the eager one bit for bit over several steps (the test the default mode cannot pass: fp32 atomics)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.ops import native

assert native.deterministic()
torch.manual_seed(0)
B, H, W = 2, 160, 224
model = RetinaNet(80, "resnet50")
g = torch.Generator().manual_seed(3)
img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g)
gt = torch.tensor([[[30., 40., 150., 140., 3.], [60., 20., 200., 120., 7.]]] * B)
cnt = torch.full((B,), 2, dtype=torch.int32)
hw = torch.tensor([[H, W]] * B, dtype=torch.int32)


def run(use_graph):
    nt = NativeTrainer(model=model, device="cuda", batch=B, image_hw=(H, W), lr=1e-4, clipnorm=0.001, use_graph=use_graph)
    nt.load_from_model(model)
    losses = []
    for _ in range(6):
        losses.append(nt.train_step(img, gt, cnt, hw).clone())
    torch.cuda.synchronize()
    return nt.P.clone(), torch.stack(losses)


pa, la = run(False)
pb, lb = run(False)
assert torch.equal(pa, pb) and torch.equal(la, lb), "two eager runs differ in deterministic mode"
pg, lg = run(True)          # steps 3.. are graph replays
assert torch.equal(pa, pg) and torch.equal(la, lg), "graph replay differs from eager in deterministic mode"
print("DETERMINISTIC_OK", la[:, 0].tolist())
