"""`--multi-gpu 2`: two native towers in ONE process tied by the P2P allreduce over peer pointers."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from batchai_retinanet_horovod_coco_b200.engine.multi_tower import MultiTowerTrainer
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet

torch.manual_seed(0)
B, H, W, K = 4, 128, 192, 4
g = torch.Generator().manual_seed(1)
img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g)
xy = torch.rand(B, 3, 2, generator=g) * torch.tensor([W * 0.5, H * 0.5])
wh = torch.rand(B, 3, 2, generator=g) * torch.tensor([W * 0.4, H * 0.4]) + 16
gt = torch.cat([xy, xy + wh, torch.randint(0, K, (B, 3, 1), generator=g).float()], dim=2)
cnt = torch.full((B,), 3, dtype=torch.int32)
hw = torch.tensor([[H, W]] * B, dtype=torch.int32)

mt = MultiTowerTrainer(RetinaNet(K, "resnet50"), 2, lr=1e-4, clipnorm=0.0, batch=B, image_hw=(H, W), shape_bucket=1)
assert torch.equal(mt.towers[0].P.cpu(), mt.towers[1].P.cpu()), "towers must start from tower 0's weights"
losses = [mt.train_step(img, gt, cnt, hw)[0].item() for _ in range(8)]
assert all(l == l for l in losses) and losses[-1] < losses[0], losses
for t in mt.towers:
    torch.cuda.synchronize(t.device)
assert torch.equal(mt.towers[0].P.cpu(), mt.towers[1].P.cpu()), "towers diverged"
assert mt.iterations == 8
# the two towers saw different halves of the batch, so their local gradients differ but the applied one is shared
assert not torch.equal(mt.towers[0].G.cpu(), mt.towers[1].G.cpu())
assert torch.equal(mt.towers[0].Gavg.cpu(), mt.towers[1].Gavg.cpu())
b, s, l = mt.predict(img[:1])
assert b.shape == (1, 300, 4)
mt.close()
print("TOWERS_OK", losses[0], losses[-1])
