"""End-to-end images/sec from JPEG files on disk (generated COCO-layout set) through the real input path:
decode on worker threads -> pinned upload -> GPU warp + resize + pad -> native training step; next to the same loop fed
from preassembled pinned batches (no decode, no augmentation).  Prints one JSON line."""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cv2
import numpy as np
import torch
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer
from batchai_retinanet_horovod_coco_b200.preprocessing.coco import CocoGenerator
from batchai_retinanet_horovod_coco_b200.preprocessing.gpu_batch import GpuBatchAssembler
from batchai_retinanet_horovod_coco_b200.preprocessing.prefetch import BatchPrefetcher
from batchai_retinanet_horovod_coco_b200.utils.transform import random_transform_generator

B = int(os.environ.get("BATCH", 8))
steps = int(os.environ.get("STEPS", 40))
workers = int(os.environ.get("WORKERS", min(32, os.cpu_count() or 1)))
n_img = int(os.environ.get("IMAGES", 192))
root = tempfile.mkdtemp(prefix="b200_coco_")
os.makedirs(f"{root}/annotations"); os.makedirs(f"{root}/images/train2017")
rng = np.random.default_rng(0)
images, anns = [], []
for i in range(n_img):
    h, w = (480, 640) if i % 3 else (640, 480)
    img = cv2.GaussianBlur(rng.integers(0, 256, (h // 8, w // 8, 3), dtype=np.uint8), (5, 5), 1.5)
    img = cv2.resize(img, (w, h), interpolation=cv2.INTER_CUBIC)
    cv2.imwrite(f"{root}/images/train2017/{i:06d}.jpg", img, [cv2.IMWRITE_JPEG_QUALITY, 90])
    images.append({"id": i, "file_name": f"{i:06d}.jpg", "height": h, "width": w})
    for k in range(6):
        x, y = rng.uniform(0, w * 0.6), rng.uniform(0, h * 0.6)
        anns.append({"id": len(anns), "image_id": i, "category_id": 1 + int(rng.integers(0, 80)), "iscrowd": 0,
                     "bbox": [float(x), float(y), float(rng.uniform(30, w * 0.35)), float(rng.uniform(30, h * 0.35))], "area": 1.0})
cats = [{"id": c + 1, "name": f"c{c}"} for c in range(80)]
json.dump({"images": images, "annotations": anns, "categories": cats}, open(f"{root}/annotations/instances_train2017.json", "w"))

dev = torch.device("cuda", 0)
tg = random_transform_generator(flip_x_chance=0.5)      # the reference's default augmentation (train.py:193)
gen = CocoGenerator(root, "train2017", transform_generator=tg, batch_size=B, image_min_side=800, image_max_side=1333, seed=1)
nt = NativeTrainer("resnet50", 80, device=dev, batch=B, image_hw=(800, 1067), shape_bucket=32, max_plans=3)
out = {"batch": B, "steps": steps, "workers": workers, "cpu_count": os.cpu_count(), "images_on_disk": n_img}


def loop(batches, n):
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(n):
        nt.train_step_async(*batches.next()).loss_cpu()
    torch.cuda.synchronize()
    return B * n / (time.time() - t0)


for mode in ("gpu_augment", "host_augment"):
    asm = GpuBatchAssembler(dev) if mode == "gpu_augment" else None
    with BatchPrefetcher(gen, workers=workers, max_queue_size=16, pin=True, assembler=asm) as bp:
        loop(bp, 12)                   # warm-up: plans, graphs, page cache
        out[mode + "_img_s"] = round(loop(bp, steps), 1)


class Fixed:
    def __init__(self, items):
        self.items, self.i = items, 0

    def next(self):
        self.i += 1
        return self.items[self.i % len(self.items)]


fixed = Fixed([tuple(t.pin_memory() for t in gen.next_compact()) for _ in range(4)])
loop(fixed, 8)
out["preassembled_img_s"] = round(loop(fixed, steps), 1)
out["gpu_augment_fraction_of_preassembled"] = round(out["gpu_augment_img_s"] / out["preassembled_img_s"], 3)
print(json.dumps(out))
