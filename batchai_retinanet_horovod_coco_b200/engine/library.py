"""Library baseline trainer: the same RetinaNet in PyTorch eager with cuDNN/cuBLAS kernels,
``torchvision`` focal loss and NCCL (DistributedDataParallel) gradient allreduce.

This is "the baseline, not the product" (SURVEY.md 6.2): the strongest
honest proxy for the reference's Horovod-NCCL + cuDNN stack on a B200, since the reference
itself (``/root/reference/train.py``: TF1/Keras 2.1/Horovod) cannot be installed offline.
It also runs on CPU (gloo) and is the oracle for the native engine's end-to-end tests.
"""

from __future__ import annotations

import torch

from ..models.retinanet import RetinaNet
from ..ops import anchors as anchor_ops
from ..ops import losses as loss_ops

CAFFE_MEAN_BGR = (103.939, 116.779, 123.68)


def assign_targets_batched(anchors: torch.Tensor, gt: torch.Tensor, gt_count: torch.Tensor, img_hw: torch.Tensor,
                           neg_thr: float = 0.4, pos_thr: float = 0.5):
    """Vectorised torch version of ops.anchors.assign_targets for a padded batch.

    anchors (A,4); gt (B,G,5); gt_count (B,); img_hw (B,2) -> state (B,A) int8, cls (B,A) int64,
    reg_target (B,A,4) fp32.
    """
    B, G, _ = gt.shape
    A = anchors.shape[0]
    a = anchors.view(1, A, 1, 4).double()
    g = gt[..., :4].view(B, 1, G, 4).double()
    iw = (torch.minimum(a[..., 2], g[..., 2]) - torch.maximum(a[..., 0], g[..., 0]) + 1).clamp(min=0)
    ih = (torch.minimum(a[..., 3], g[..., 3]) - torch.maximum(a[..., 1], g[..., 1]) + 1).clamp(min=0)
    inter = iw * ih
    ua = ((a[..., 2] - a[..., 0] + 1) * (a[..., 3] - a[..., 1] + 1) +
          (g[..., 2] - g[..., 0] + 1) * (g[..., 3] - g[..., 1] + 1) - inter).clamp(min=1e-8)
    iou = inter / ua
    valid = torch.arange(G, device=gt.device).view(1, 1, G) < gt_count.view(B, 1, 1)
    iou = torch.where(valid, iou, torch.full_like(iou, -1.0))
    best, idx = iou.max(dim=2)
    # torch.where throughout (no boolean-mask indexing): no host sync, so the step can be captured in a CUDA graph
    m1 = torch.full((B, A), -1, dtype=torch.int8, device=gt.device)
    state = torch.where(best < neg_thr, torch.zeros_like(m1), m1)
    state = torch.where(best >= pos_thr, torch.ones_like(m1), state)
    state = torch.where((gt_count == 0).view(B, 1), torch.zeros_like(m1), state)
    cx = (anchors[:, 0] + anchors[:, 2]) / 2
    cy = (anchors[:, 1] + anchors[:, 3]) / 2
    outside = (cx.view(1, A) >= img_hw[:, 1].view(B, 1)) | (cy.view(1, A) >= img_hw[:, 0].view(B, 1))
    state = torch.where(outside, m1, state)
    matched = torch.gather(gt, 1, idx.unsqueeze(-1).expand(B, A, 5))
    aw = (anchors[:, 2] - anchors[:, 0]).view(1, A)
    ah = (anchors[:, 3] - anchors[:, 1]).view(1, A)
    reg = torch.stack([(matched[..., 0] - anchors[:, 0]) / aw, (matched[..., 1] - anchors[:, 1]) / ah,
                       (matched[..., 2] - anchors[:, 2]) / aw, (matched[..., 3] - anchors[:, 3]) / ah], dim=-1) / 0.2
    return state, matched[..., 4].long(), reg.float()


class LibraryTrainer:
    """One training step = H2D of uint8 images + boxes, forward, losses, backward, clip, allreduce, Adam."""

    def __init__(self, backbone: str = "resnet50", num_classes: int = 80, device="cuda", dtype=torch.bfloat16,
                 lr: float = 1e-5, clipnorm: float = 0.001, distributed: bool = False, model: RetinaNet | None = None,
                 freeze_backbone: bool = False, use_graph: bool | None = None):
        self.device = torch.device(device)
        # single-process CUDA runs replay the whole step (forward, losses, backward, clip, fused Adam) as ONE CUDA graph
        # after three eager steps per input shape: the baseline should not lose to launch overhead what it does not
        # lose to its kernels
        if use_graph is None:
            use_graph = torch.device(device).type == "cuda" and not distributed
        self.use_graph = bool(use_graph)
        self._graphs = {}
        self.dtype = dtype
        self.num_classes = num_classes
        self.clipnorm = clipnorm
        self.model = (model or RetinaNet(num_classes, backbone)).to(self.device)
        self.num_classes = self.model.num_classes
        if freeze_backbone:
            self.model.freeze_backbone()
        if self.device.type == "cuda":
            self.model = self.model.to(memory_format=torch.channels_last)
        self.train_model = self.model
        if distributed:
            from torch.nn.parallel import DistributedDataParallel as DDP
            ids = [self.device.index] if self.device.type == "cuda" else None
            self.train_model = DDP(self.model, device_ids=ids, bucket_cap_mb=64, gradient_as_bucket_view=True)
        params = [p for p in self.model.parameters() if p.requires_grad]
        kw = {"fused": True, "capturable": True} if self.device.type == "cuda" else {}
        self.opt = torch.optim.Adam(params, lr=lr, betas=(0.9, 0.999), eps=1e-7, **kw)
        self.mean = torch.tensor(CAFFE_MEAN_BGR, device=self.device).view(1, 3, 1, 1)
        self._anchors = {}

    def anchors(self, hw, feature_shapes=None):
        if hw not in self._anchors:
            self._anchors[hw] = anchor_ops.anchors_for_shape(hw, feature_shapes=feature_shapes).to(self.device)
        return self._anchors[hw]

    def loss(self, images_u8: torch.Tensor, gt: torch.Tensor, gt_count: torch.Tensor, img_hw: torch.Tensor):
        x = images_u8.permute(0, 3, 1, 2).float()
        x = x / 127.5 - 1.0 if getattr(self.model, "preprocess_mode", "caffe") == "tf" else x - self.mean
        if self.device.type == "cuda":
            x = x.contiguous(memory_format=torch.channels_last)
        with torch.autocast(self.device.type, dtype=self.dtype, enabled=self.dtype != torch.float32):
            out = self.train_model(x)
        h, w = images_u8.shape[1:3]
        anchors = self.anchors((h, w), out["feature_shapes"])
        with torch.no_grad():
            state, cls, reg_t = assign_targets_batched(anchors, gt, gt_count, img_hw)
        npos = (state == 1).sum()
        lc = loss_ops.focal_compact(out["classification"], state, cls, num_pos=npos)
        lr_ = loss_ops.smooth_l1_compact(out["regression"], reg_t, state, num_pos=npos)
        return lc + lr_, lc, lr_

    def train_step(self, images_u8, gt, gt_count, img_hw):
        if self.use_graph:
            return self._graph_step(images_u8, gt, gt_count, img_hw)
        return self._eager_step(images_u8, gt, gt_count, img_hw, set_to_none=True)

    def _graph_step(self, images_u8, gt, gt_count, img_hw):
        key = (tuple(images_u8.shape), tuple(gt.shape))
        st = self._graphs.get(key)
        if st is None:
            st = self._graphs[key] = {"warm": 0}
        if "graph" not in st:
            st["warm"] += 1
            if st["warm"] <= 3:
                return self._eager_step(images_u8, gt, gt_count, img_hw, set_to_none=False)
            dev = self.device
            st["in"] = [torch.empty_like(t, device=dev) for t in (images_u8, gt, gt_count, img_hw)]
            for d, t in zip(st["in"], (images_u8, gt, gt_count, img_hw)):
                d.copy_(t, non_blocking=True)
            torch.cuda.synchronize(dev)
            g = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(g):
                    st["out"] = self._eager_step(*st["in"], set_to_none=False)
                st["graph"] = g
            except Exception as e:  # noqa: BLE001 - capture is an optimisation of the baseline
                import warnings
                warnings.warn(f"library baseline: CUDA-graph capture failed ({e}); staying eager")
                self.use_graph = False
                return self._eager_step(images_u8, gt, gt_count, img_hw, set_to_none=True)
        for d, t in zip(st["in"], (images_u8, gt, gt_count, img_hw)):
            d.copy_(t, non_blocking=True)
        st["graph"].replay()
        return st["out"].clone()

    def _eager_step(self, images_u8, gt, gt_count, img_hw, set_to_none=True):
        dev = self.device
        images_u8 = images_u8.to(dev, non_blocking=True)
        gt = gt.to(dev, non_blocking=True)
        gt_count = gt_count.to(dev, non_blocking=True)
        img_hw = img_hw.to(dev, non_blocking=True)
        self.opt.zero_grad(set_to_none=set_to_none)
        total, lc, lr_ = self.loss(images_u8, gt, gt_count, img_hw)
        total.backward()
        if self.clipnorm:
            torch.nn.utils.clip_grad_norm_([p for p in self.model.parameters() if p.requires_grad], self.clipnorm,
                                           foreach=True if self.device.type == "cuda" else None)
        self.opt.step()
        return torch.stack([total.detach(), lr_.detach(), lc.detach()])
