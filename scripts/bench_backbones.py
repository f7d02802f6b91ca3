"""Non-ResNet backbones (reference train.py:323-324 `--backbone`): the native engine of each family (VGGTrainer,
MobileNetTrainer, DenseNetTrainer) vs the library baseline (CUDA-graphed), one GPU, 800x1333, batch 8, device-timed."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench_configs_common import batch, timeit, dev
from batchai_retinanet_horovod_coco_b200.engine import native_engine_class
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.engine.library import LibraryTrainer

B, H, W = 8, 800, 1333
for bb in sys.argv[1:] or ("vgg16", "mobilenet224_1.0", "densenet121"):
    for impl in ("b200", "library"):
        torch.cuda.empty_cache()
        b = batch(B, H, W)
        if impl == "library":
            tr = LibraryTrainer(bb, 80, device=dev)
        else:
            model = RetinaNet(80, bb)
            tr = native_engine_class(model)(model=model, device=dev, batch=B, image_hw=(H, W))
            impl = "b200-" + type(tr).__name__
        ms = timeit(lambda: tr.train_step(*b), 10, warm=6)          # graph capture (4th call of the library arm) stays out of the timed region
        row = {"backbone": bb, "impl": impl, "batch": B, "image": [H, W], "train_ms_per_step": round(ms, 3),
               "train_img_per_s": round(B / ms * 1e3, 1)}
        if impl != "library":
            row["graph"] = bool(tr.use_graph)
            row["phases_ms"] = {k: round(v, 2) for k, v in tr.profile_step(*b).items()}
        print(json.dumps(row), flush=True)
        del tr
