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
ref2 = copy.deepcopy(model).cuda()
img, gt, cnt, hw = _inputs(B, H, W)
ht = create_trainer(model, "cuda", batch=B, image_hw=(H, W))
loss_h, pl = ht.forward_loss(img, gt, cnt, hw)
torch.cuda.synchronize()
Ps = [t.clone() for t in pl["fpn"]["P"]]
ht._infer = True
with torch.no_grad():
    c3, c4, c5 = ht._forward_backbone(pl, img.cuda())
    fp2 = ref2.fpn(c3.permute(0, 3, 1, 2).float(), c4.permute(0, 3, 1, 2).float(), c5.permute(0, 3, 1, 2).float())
    for a, b in zip(Ps, fp2):
        print("P vs plain-copy ref", tuple(a.shape), _cos(a.float(), b.permute(0, 2, 3, 1)))
    lib = LibraryTrainer(model=ref_model, device="cuda", dtype=torch.float32)
    fp3 = ref_model.fpn(c3.permute(0, 3, 1, 2).float(), c4.permute(0, 3, 1, 2).float(), c5.permute(0, 3, 1, 2).float())
    for a, b in zip(Ps, fp3):
        print("P vs LibraryTrainer-model ref", tuple(a.shape), _cos(a.float(), b.permute(0, 2, 3, 1)), b.shape, b.stride())
    print("w equal", torch.equal(ref_model.fpn.P5.weight, ref2.fpn.P5.weight), ref_model.fpn.P5.weight.stride(), ref2.fpn.P5.weight.stride())
