import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer
from batchai_retinanet_horovod_coco_b200 import models
from batchai_retinanet_horovod_coco_b200.preprocessing.synthetic import SyntheticGenerator
torch.manual_seed(0)
gen = SyntheticGenerator(num_images=8, image_hw=(256, 320), classes=3, boxes_per_image=3, seed=3, batch_size=8,
                         image_min_side=256, image_max_side=320, shuffle_groups=False)
model = models.backbone("resnet50").retinanet(3)
tr = NativeTrainer(model=model, device="cuda", lr=1e-4, clipnorm=0.0, batch=8, image_hw=(256, 320), use_graph=False)
w0 = tr.cls_out.wview(tr.P).clone(); t0 = tr.cls_tower[0].wview(tr.P).clone(); r0 = tr.reg_out.wview(tr.P).clone()
for s in range(30):
    batch = gen.next_compact()
    loss = tr.train_step(*batch)
    if s in (0, 1, 5, 29):
        p = tr._plan(8, 256, 320)
        lo = torch.cat([t.float().flatten() for t in p["cls_o"]])
        c5 = p["C"][2].float()
        g = tr.gradients_by_layer()
        print(s, "loss", [round(v, 3) for v in loss.tolist()], "logits absmax %.3g mean %.3g" % (lo.abs().max(), lo.mean()),
              "C5 absmax %.3g" % c5.abs().max(), "P3 absmax %.3g" % p["fpn"]["P"][0].float().abs().max(),
              "tower3 absmax %.3g" % p["cls_t"][3][0].float().abs().max(),
              "|g cls_out| %.3g |g cls_t0| %.3g |g reg_out| %.3g" % (g["pyramid_classification"][0].norm(), g["pyramid_classification_0"][0].norm(), g["pyramid_regression"][0].norm()),
              "dcls absmax %.3g" % torch.cat([t.float().flatten() for t in p["dcls_o"]]).abs().max(),
              "scal", tr.scal.tolist()[:4], "npos", tr.npos.item())
print("dW cls_out %.3g  cls_t0 %.3g  reg_out %.3g" % ((tr.cls_out.wview(tr.P) - w0).abs().max(), (tr.cls_tower[0].wview(tr.P) - t0).abs().max(), (tr.reg_out.wview(tr.P) - r0).abs().max()))
