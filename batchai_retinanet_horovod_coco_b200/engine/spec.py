"""Spec-layer trainer: the PyTorch ``RetinaNet`` module + autograd + ``KerasAdam`` wrapped in the
Horovod-style ``DistributedOptimizer``.  Runs anywhere (CPU included) and is what the reference's
"``mpiexec -n 2 python train.py ...``" plumbing configuration maps to when no GPU is present
(``train.py:18``; BASELINE.json config 1).  Same public surface as ``NativeTrainer``.
"""

from __future__ import annotations

import torch

from .. import parallel as hvd
from ..models.retinanet import RetinaNet, retinanet_bbox
from ..ops import anchors as anchor_ops
from ..ops import losses as loss_ops
from ..parallel.optimizer import DistributedOptimizer, KerasAdam
from ..utils import keras_h5
from .library import CAFFE_MEAN_BGR, assign_targets_batched


class SpecTrainer:
    def __init__(self, model: RetinaNet, device="cpu", lr: float = 1e-5, clipnorm: float = 0.001,
                 distributed: bool | None = None):
        self.device = torch.device(device)
        self.model = model.to(self.device)
        self.num_classes = model.num_classes
        params = [p for p in self.model.parameters() if p.requires_grad]
        self.base_optimizer = KerasAdam(params, lr=lr, clipnorm=clipnorm)
        if distributed is None:
            distributed = hvd.is_initialized() and hvd.size() > 1
        named = [(n, p) for n, p in self.model.named_parameters() if p.requires_grad]
        self.optimizer = DistributedOptimizer(self.base_optimizer, named_parameters=named) if distributed \
            else self.base_optimizer
        self.mean = torch.tensor(CAFFE_MEAN_BGR, device=self.device).view(1, 3, 1, 1)
        self.prediction_model = retinanet_bbox(self.model)
        self._anchors = {}

    # ---- hyper-parameters ---------------------------------------------------------------------
    @property
    def lr(self):
        return self.base_optimizer.param_groups[0]["lr"]

    @lr.setter
    def lr(self, v):
        for g in self.base_optimizer.param_groups:
            g["lr"] = float(v)

    @property
    def iterations(self):
        return self.base_optimizer.iterations

    # ---- steps --------------------------------------------------------------------------------
    def _prep(self, images_u8):
        x = images_u8.to(self.device).permute(0, 3, 1, 2).float()
        if getattr(self.model, "preprocess_mode", "caffe") == "tf":
            return x / 127.5 - 1.0      # keras-retinanet preprocess_image(mode='tf'): mobilenet / densenet
        return x - self.mean

    def train_step(self, images_u8, gt, gt_count, img_hw):
        dev = self.device
        self.model.train()
        x = self._prep(images_u8)
        out = self.model(x)
        hw = tuple(images_u8.shape[1:3])
        if hw not in self._anchors:
            # pyramid shapes come from the tensors (VGG / DenseNet pools floor, ResNet ceils)
            self._anchors[hw] = anchor_ops.anchors_for_shape(hw, feature_shapes=out["feature_shapes"]).to(dev)
        with torch.no_grad():
            state, cls, reg_t = assign_targets_batched(self._anchors[hw], gt.to(dev).float(), gt_count.to(dev),
                                                       img_hw.to(dev))
        npos = (state == 1).sum()
        lc = loss_ops.focal_compact(out["classification"], state, cls, num_pos=npos)
        lr_ = loss_ops.smooth_l1_compact(out["regression"], reg_t, state, num_pos=npos)
        self.base_optimizer.zero_grad(set_to_none=False)
        (lc + lr_).backward()
        self.optimizer.step()
        return torch.stack([(lc + lr_).detach(), lr_.detach(), lc.detach()])

    @torch.no_grad()
    def predict(self, images_u8, **kw):
        self.model.eval()
        return self.prediction_model(self._prep(images_u8))

    # ---- state --------------------------------------------------------------------------------
    def broadcast_state(self, root: int = 0):
        if hvd.is_initialized() and hvd.size() > 1:
            hvd.broadcast_parameters(self.model.state_dict(), root)
            hvd.broadcast_optimizer_state(self.base_optimizer, root)

    def store_to_model(self):
        return self.model

    def optimizer_slots(self):
        out = {}
        for name, t in keras_h5.trainable_weight_list(self.model):
            st = self.base_optimizer.state.get(t)
            if st:
                out[name] = (st["m"], st["v"])
        return out

    def load_optimizer_state(self, state: dict):
        cfg = state.get("optimizer_config") or {}
        if "lr" in cfg:
            self.lr = cfg["lr"]
        if "clipnorm" in cfg:
            self.base_optimizer.clipnorm = cfg["clipnorm"]
        self.base_optimizer.iterations = int(state.get("iterations", 0))
        for name, t in keras_h5.trainable_weight_list(self.model):
            if name in state.get("slots", {}):
                m, v = state["slots"][name]
                self.base_optimizer.state[t] = {"m": m.to(t.device, t.dtype).clone(), "v": v.to(t.device, t.dtype).clone()}

    def save(self, path: str, epoch: int | None = None, weights_only: bool = False):
        if weights_only:
            keras_h5.save_weights(self.model, path)
        else:
            keras_h5.save_model(self.model, path, optimizer_config=self.base_optimizer.get_config(),
                                optimizer_slots=self.optimizer_slots(), iterations=self.iterations, epoch=epoch)
