"""``--multi-gpu N``: single-process data parallelism over N GPUs.

The reference inherits this mode from upstream (``keras.utils.multi_gpu_model``: model on ``/cpu:0``,
one tower per GPU, batch split, ``train.py:31,86-89,310-321,367-368``).  It is
legacy next to the one-process-per-GPU path, but a user of the reference can ask for it.

Here a tower is a full ``NativeTrainer`` on its own GPU driven by its own host thread (kernel launches
release the GIL), and the towers are tied together by the same fused P2P allreduce kernel as the
multi-process path -- the regions are shared as raw peer pointers instead of CUDA IPC handles
(``parallel/p2p.py:LocalP2PComm``).  Differences from ``multi_gpu_model``, on purpose:

* weights live on every GPU (kept bit-identical by the allreduce) instead of on the CPU, so there is no
  per-step weight download and no gradient upload;
* each tower normalises its losses by its own positive-anchor count and the gradients are averaged,
  i.e. the Horovod semantics of the main path, not one normaliser over the concatenated batch.
"""

from __future__ import annotations

import copy
from concurrent.futures import ThreadPoolExecutor

import torch

from ..parallel.p2p import LocalGroup, LocalP2PComm


class MultiTowerTrainer:
    def __init__(self, model, num_towers: int, lr=1e-5, clipnorm=0.001, batch: int = 2, image_hw=(800, 1333),
                 freeze_backbone: bool = False, shape_bucket: int = 64):
        if not torch.cuda.is_available() or torch.cuda.device_count() < num_towers:
            raise RuntimeError(f"--multi-gpu {num_towers} needs {num_towers} visible GPUs, "
                               f"found {torch.cuda.device_count() if torch.cuda.is_available() else 0}")
        if batch < num_towers:
            raise ValueError(f"Batch size ({batch}) must be equal to or higher than the number of GPUs ({num_towers})")
        from . import native_engine_class
        cls = native_engine_class(model)
        self.n = num_towers
        self.model = model
        self.device = torch.device("cuda", 0)     # where losses are gathered and predictions run
        self._pool = ThreadPoolExecutor(max_workers=num_towers, thread_name_prefix="tower")
        group = LocalGroup(num_towers)
        models = [model] + [copy.deepcopy(model) for _ in range(num_towers - 1)]

        self._group = group

        def build(i):
            dev = torch.device("cuda", i)
            torch.cuda.set_device(dev)
            comm = LocalP2PComm(i, num_towers, dev, group)
            return cls(model=models[i], device=dev, lr=lr, clipnorm=clipnorm, batch=-(-batch // num_towers),
                       image_hw=image_hw, freeze_backbone=freeze_backbone, comm=comm, shape_bucket=shape_bucket)

        self.towers = self._all(build)

    def _all(self, fn):
        """Run ``fn(i)`` on every tower's thread; a tower that raises releases the others from the barrier."""
        def guarded(i):
            try:
                return fn(i)
            except BaseException:
                self._group.abort()
                raise
        return list(self._pool.map(guarded, range(self.n)))

    # ---- hyper-parameters (kept equal on every tower) ------------------------------------------
    @property
    def lr(self):
        return self.towers[0].lr

    @lr.setter
    def lr(self, v):
        for t in self.towers:
            t.lr = float(v)

    @property
    def iterations(self):
        return self.towers[0].iterations

    @property
    def num_classes(self):
        return self.towers[0].num_classes

    # ---- steps ---------------------------------------------------------------------------------
    def _split(self, t):
        """Split the batch axis into N contiguous chunks (``multi_gpu_model`` slices the same way)."""
        # every tower gets floor(B / n) samples and the last one also the remainder, as keras.utils.multi_gpu_model
        # does (reference train.py:86-89); ceil-sized chunks would leave the last tower empty for e.g. B=5, n=4
        B = t.shape[0]
        per = B // self.n
        return [t[i * per:(B if i == self.n - 1 else (i + 1) * per)] for i in range(self.n)]

    def train_step(self, images_u8, gt, gt_count, img_hw):
        if images_u8.shape[0] < self.n:
            raise ValueError(f"batch of {images_u8.shape[0]} cannot be split over {self.n} towers")
        parts = list(zip(self._split(images_u8), self._split(gt), self._split(gt_count), self._split(img_hw)))

        def run(i):
            torch.cuda.set_device(i)
            return self.towers[i].train_step(*parts[i])

        losses = self._all(run)
        w = torch.tensor([p[0].shape[0] for p in parts], dtype=torch.float32)
        stacked = torch.stack([l.to("cuda:0") for l in losses])        # [N, 3]
        return (stacked * (w / w.sum()).to("cuda:0").view(-1, 1)).sum(dim=0)

    def predict(self, images_u8, **kw):
        with torch.cuda.device(0):
            return self.towers[0].predict(images_u8, **kw)

    # ---- state ---------------------------------------------------------------------------------
    def broadcast_state(self, root: int = 0):
        def run(i):
            torch.cuda.set_device(i)
            self.towers[i].broadcast_state(root)
        self._all(run)

    def store_to_model(self):
        return self.towers[0].store_to_model()

    def optimizer_config(self):
        return self.towers[0].optimizer_config()

    def optimizer_slots(self):
        return self.towers[0].optimizer_slots()

    def load_optimizer_state(self, state: dict):
        for t in self.towers:
            t.load_optimizer_state(state)

    def save(self, path: str, epoch=None, weights_only: bool = False):
        with torch.cuda.device(0):
            return self.towers[0].save(path, epoch=epoch, weights_only=weights_only)

    def close(self):
        self._pool.shutdown(wait=True)
