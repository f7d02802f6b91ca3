import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_gpu_engine import _cos, _inputs
from test_gpu_hybrid import _model
from batchai_retinanet_horovod_coco_b200.engine import create_trainer
from batchai_retinanet_horovod_coco_b200.engine.library import LibraryTrainer
name = sys.argv[1] if len(sys.argv) > 1 else "vgg16"
B, H, W = 2, 160, 224
model = _model(name)
ref_model = copy.deepcopy(model)
img, gt, cnt, hw = _inputs(B, H, W)
ht = create_trainer(model, "cuda", batch=B, image_hw=(H, W))
loss_h, p = ht.forward_loss(img, gt, cnt, hw)
pl = ht._plan(B, H, W)
lib = LibraryTrainer(model=ref_model, device="cuda", dtype=torch.float32)
ref_model.train()
x = img.cuda().permute(0, 3, 1, 2).float()
x = x / 127.5 - 1.0 if ref_model.preprocess_mode == "tf" else x - lib.mean
feats = ref_model.backbone(x)[1:]
ht._infer = False
c = ht._forward_backbone(pl, img.cuda())
for a, b in zip(c, feats):
    b = b.permute(0, 2, 3, 1)
    print("feat", tuple(a.shape), "cos", _cos(a.float(), b), "norm ratio", float(a.float().norm() / b.norm()), "absmax", float(b.abs().max()))
fp = ref_model.fpn(*feats)
for a, b in zip(pl["fpn"]["P"], fp):
    print("P", tuple(a.shape), _cos(a.float(), b.permute(0, 2, 3, 1)))
print("loss", loss_h.tolist())
total, lc, lr_ = lib.loss(img.cuda(), gt.cuda(), cnt.cuda(), hw.cuda())
print("ref", total.item(), lr_.item(), lc.item())
