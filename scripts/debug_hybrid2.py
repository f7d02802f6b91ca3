import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import torch.nn.functional as F
from test_gpu_engine import _cos, _inputs
from test_gpu_hybrid import _model
from batchai_retinanet_horovod_coco_b200.engine import create_trainer
name = sys.argv[1] if len(sys.argv) > 1 else "vgg16"
B, H, W = 2, 160, 224
model = _model(name)
ref_model = copy.deepcopy(model).cuda()
img, gt, cnt, hw = _inputs(B, H, W)
ht = create_trainer(model, "cuda", batch=B, image_hw=(H, W))
pl = ht._plan(B, H, W)
with torch.no_grad():
    ht._infer = True
    ht._forward(pl, img.cuda())
    torch.cuda.synchronize()
    c3, c4, c5 = pl["C"]
    f = ref_model.fpn
    c5r = f.C5_reduced(c5.permute(0, 3, 1, 2).float())
    print("c5 contiguous", c5.is_contiguous(), c5.shape, c5.stride(), c5.dtype)
    print("c5r cos", _cos(pl["fpn"]["c5r"].float(), c5r.permute(0, 2, 3, 1)))
    print("weights equal", torch.equal(model.fpn.C5_reduced.weight.cpu(), ref_model.fpn.C5_reduced.weight.cpu()))
    wv = ht.fpn["C5_reduced"].wview(ht.P)
    print("P master vs module", _cos(wv.view(1, 1, 256, -1).permute(2, 3, 0, 1), ref_model.fpn.C5_reduced.weight))
    print("wf vs module", _cos(ht.fpn["C5_reduced"].wf.float()[0, :, :c5.shape[-1]], ref_model.fpn.C5_reduced.weight[:, :, 0, 0]))
    print("n_train", ht.n_train, "n_params", ht.n_params, "C5_reduced off", ht.fpn["C5_reduced"].w_off, "extra0", ht._extra[0][1:], "last", ht._extra[-1][1:])
    fp = f(c3.permute(0, 3, 1, 2).float(), c4.permute(0, 3, 1, 2).float(), c5.permute(0, 3, 1, 2).float())
    for a, b in zip(pl["fpn"]["P"], fp):
        print("infer P", tuple(a.shape), _cos(a.float(), b.permute(0, 2, 3, 1)), float(a.float().abs().max()), float(b.abs().max()))
ht._infer = False
loss_h, _ = ht.forward_loss(img, gt, cnt, hw)
torch.cuda.synchronize()
for a, b in zip(pl["fpn"]["P"], fp):
    print("after fwd/bwd P", tuple(a.shape), _cos(a.float(), b.permute(0, 2, 3, 1)))
