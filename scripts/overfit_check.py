"""End-to-end learning check: overfit a tiny synthetic set with the native engine, then run the prediction
graph + VOC-style evaluation on the same images.  Exercises target assignment, both losses, backward, Adam,
box decode, NMS and the evaluator together: mAP must climb well above chance."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.engine import create_trainer
from batchai_retinanet_horovod_coco_b200 import models
from batchai_retinanet_horovod_coco_b200.preprocessing.synthetic import SyntheticGenerator
from batchai_retinanet_horovod_coco_b200.utils.eval import evaluate, mean_ap

steps = int(os.environ.get("STEPS", 600))
bb = os.environ.get("BACKBONE", "resnet50")
torch.manual_seed(0)
gen = SyntheticGenerator(num_images=8, image_hw=(256, 320), classes=3, boxes_per_image=3, seed=3, batch_size=8,
                         image_min_side=256, image_max_side=320, shuffle_groups=False)
model = models.backbone(bb).retinanet(gen.num_classes())
tr = create_trainer(model, "cuda", lr=float(os.environ.get("LR", 1e-4)), clipnorm=0.0, batch=8, image_hw=(256, 320))
t0 = time.time()
log = []
for s in range(steps):
    loss = tr.train_step(*gen.next_compact(pin=True))
    if s % 100 == 0 or s == steps - 1:
        log.append((s, [round(v, 4) for v in loss.tolist()]))
torch.cuda.synchronize()
aps = evaluate(gen, lambda x: tr.predict(x), iou_threshold=0.5, score_threshold=0.05, max_detections=100)
print(json.dumps({"backbone": bb, "steps": steps, "train_s": round(time.time() - t0, 1), "loss_log": log,
                  "mAP@0.5": round(mean_ap(aps), 4), "per_class": {k: [round(v[0], 3), v[1]] for k, v in aps.items()}}))
