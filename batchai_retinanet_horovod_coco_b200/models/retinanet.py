"""RetinaNet training graph and prediction graph (spec layer, plain PyTorch).

Behavioural spec: keras-retinanet ``retinanet(...)`` and ``retinanet_bbox(model=...)`` as the
reference binds them at ``train.py:46,88-95,417``.

* FPN: ``C5_reduced`` 1x1 -> ``P5`` 3x3; ``P5_upsampled`` = nearest resize to C4's shape;
  ``C4_reduced`` + add -> ``P4``; same for P3; ``P6`` = 3x3/2 on **C5**; ``P7`` = 3x3/2 on
  relu(P6).  256 channels, bias, no norm, no activation on P3-P5.
* Two sub-models shared over P3..P7 (weights shared across levels): regression
  (4x 3x3 256 ReLU + 3x3 -> 9*4) and classification (4x 3x3 256 ReLU + 3x3 -> 9*K, bias
  ``-log((1-pi)/pi)``, pi = 0.01).  Kernel init N(0, 0.01).
* Outputs are concatenated over levels on the anchor axis and named ``regression`` /
  ``classification``; those are the loss-dict keys at reference ``train.py:99-102``.

The Keras model applies the sigmoid inside the classification sub-model.  Here the module
returns **logits** by default (``return_logits=True``) because the fused focal-loss kernel
works in logit space (numerically better); ``return_logits=False`` gives the
Keras behaviour (probabilities).
"""

from __future__ import annotations

import math
from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

from .resnet import ResNet, tf_same_pad


def make_backbone(name: str) -> nn.Module:
    """Backbone module by substring, like keras-retinanet's ``models.backbone`` (reference ``train.py:390``)."""
    if "resnet" in name:
        return ResNet(name)
    if "vgg" in name:
        from .vgg import VGG
        return VGG(name)
    if "mobilenet" in name:
        from .mobilenet import MobileNet
        return MobileNet(name)
    if "densenet" in name:
        from .densenet import DenseNet
        return DenseNet(name)
    raise NotImplementedError(f"Backbone class for '{name}' not implemented.")

NUM_ANCHORS = 9
FEATURE_SIZE = 256
PRIOR_PROBABILITY = 0.01


def conv_same(conv: nn.Conv2d, x: torch.Tensor) -> torch.Tensor:
    """Apply a stride-2 3x3 conv with TensorFlow 'same' padding (possibly asymmetric)."""
    k, s = conv.kernel_size[0], conv.stride[0]
    ph = tf_same_pad(x.shape[-2], k, s)
    pw = tf_same_pad(x.shape[-1], k, s)
    x = F.pad(x, (pw[0], pw[1], ph[0], ph[1]))
    return F.conv2d(x, conv.weight, conv.bias, stride=s)


def upsample_like(src: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """Keras-retinanet ``UpsampleLike``: nearest resize of ``src`` to ``target``'s HxW."""
    return F.interpolate(src, size=target.shape[-2:], mode="nearest")


class FPN(nn.Module):
    def __init__(self, c3: int, c4: int, c5: int, feature_size: int = FEATURE_SIZE):
        super().__init__()
        f = feature_size
        self.C5_reduced = nn.Conv2d(c5, f, 1)
        self.P5 = nn.Conv2d(f, f, 3, padding=1)
        self.C4_reduced = nn.Conv2d(c4, f, 1)
        self.P4 = nn.Conv2d(f, f, 3, padding=1)
        self.C3_reduced = nn.Conv2d(c3, f, 1)
        self.P3 = nn.Conv2d(f, f, 3, padding=1)
        self.P6 = nn.Conv2d(c5, f, 3, stride=2)
        self.P7 = nn.Conv2d(f, f, 3, stride=2)

    LAYER_ORDER = ["C5_reduced", "P5", "C4_reduced", "P4", "C3_reduced", "P3", "P6", "P7"]

    def keras_layers(self):
        return OrderedDict((n, getattr(self, n)) for n in self.LAYER_ORDER)

    def forward(self, c3, c4, c5):
        p5 = self.C5_reduced(c5)
        p5_up = upsample_like(p5, c4)
        p5 = self.P5(p5)
        p4 = self.C4_reduced(c4) + p5_up
        p4_up = upsample_like(p4, c3)
        p4 = self.P4(p4)
        p3 = self.C3_reduced(c3) + p4_up
        p3 = self.P3(p3)
        p6 = conv_same(self.P6, c5)
        p7 = conv_same(self.P7, F.relu(p6))
        return [p3, p4, p5, p6, p7]


class SubModel(nn.Module):
    """Shared head: 4x (3x3, 256, ReLU) then a 3x3 output conv."""

    def __init__(self, prefix: str, num_values: int, num_anchors: int = NUM_ANCHORS,
                 feature_size: int = FEATURE_SIZE, prior: float | None = None):
        super().__init__()
        self.prefix = prefix
        self.num_values = num_values
        self.num_anchors = num_anchors
        self.tower = nn.ModuleList([nn.Conv2d(feature_size, feature_size, 3, padding=1) for _ in range(4)])
        self.out = nn.Conv2d(feature_size, num_anchors * num_values, 3, padding=1)
        for m in list(self.tower) + [self.out]:
            nn.init.normal_(m.weight, mean=0.0, std=0.01)
            nn.init.zeros_(m.bias)
        if prior is not None:
            nn.init.constant_(self.out.bias, -math.log((1 - prior) / prior))

    def keras_layers(self):
        d = OrderedDict()
        for i, m in enumerate(self.tower):
            d[f"{self.prefix}_{i}"] = m
        d[self.prefix] = self.out
        return d

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for m in self.tower:
            x = F.relu(m(x))
        x = self.out(x)
        b, _, h, w = x.shape
        # Keras is NHWC: reshape (B,H,W,A*V) -> (B, H*W*A, V)
        return x.permute(0, 2, 3, 1).reshape(b, h * w * self.num_anchors, self.num_values)


class RetinaNet(nn.Module):
    """Training model.  ``forward(images NCHW) -> {'regression','classification'}``."""

    def __init__(self, num_classes: int, backbone: str = "resnet50", return_logits: bool = True):
        super().__init__()
        self.num_classes = num_classes
        self.backbone_name = backbone
        self.return_logits = return_logits
        self.backbone = make_backbone(backbone)
        self.preprocess_mode = getattr(self.backbone, "preprocess_mode", "caffe")
        c2, c3, c4, c5 = self.backbone.out_channels
        self.fpn = FPN(c3, c4, c5)
        self.regression_submodel = SubModel("pyramid_regression", 4)
        self.classification_submodel = SubModel("pyramid_classification", num_classes, prior=PRIOR_PROBABILITY)

    # -- structure helpers -------------------------------------------------------------
    def keras_layers(self) -> "OrderedDict[str, nn.Module]":
        """Ordered ``keras layer name -> module`` (flat, sub-model layers included)."""
        d = OrderedDict()
        d.update(self.backbone.keras_layers())
        d.update(self.fpn.keras_layers())
        d.update(self.regression_submodel.keras_layers())
        d.update(self.classification_submodel.keras_layers())
        return d

    def freeze_backbone(self):
        """``--freeze-backbone`` (reference ``train.py:82,375``): backbone gets no gradient."""
        for p in self.backbone.parameters():
            p.requires_grad_(False)
        return self

    def trainable_parameters(self):
        return [p for p in self.parameters() if p.requires_grad]

    def features(self, images: torch.Tensor):
        _, c3, c4, c5 = self.backbone(images)
        return self.fpn(c3, c4, c5)

    def forward(self, images: torch.Tensor):
        feats = self.features(images)
        regression = torch.cat([self.regression_submodel(f) for f in feats], dim=1)
        classification = torch.cat([self.classification_submodel(f) for f in feats], dim=1)
        if not self.return_logits:
            classification = torch.sigmoid(classification)
        return {"regression": regression, "classification": classification,
                "feature_shapes": [tuple(f.shape[-2:]) for f in feats]}

    def summary(self) -> str:
        lines = ["_" * 78, f"{'Layer (name)':<40}{'Output/param shape':<26}{'Param #':>10}", "=" * 78]
        total = trainable = 0
        for name, m in self.keras_layers().items():
            n = sum(p.numel() for p in m.parameters()) + sum(b.numel() for b in m.buffers())
            t = sum(p.numel() for p in m.parameters() if p.requires_grad)
            shape = tuple(m.weight.shape) if hasattr(m, "weight") else tuple(m.gamma.shape)
            lines.append(f"{name:<40}{str(shape):<26}{n:>10,}")
            total += n
            trainable += t
        lines += ["=" * 78, f"Total params: {total:,}", f"Trainable params: {trainable:,}",
                  f"Non-trainable params: {total - trainable:,}", "_" * 78]
        return "\n".join(lines)


class RetinaNetBBox(nn.Module):
    """Prediction graph: shares the training model's tensors (no weight copy).

    ``forward(images) -> boxes (B,300,4), scores (B,300), labels (B,300)`` padded with -1.
    Reference ``train.py:95,408`` -> ``retinanet_bbox(model=model)``.
    """

    def __init__(self, model: RetinaNet, nms: bool = True, score_threshold: float = 0.05,
                 max_detections: int = 300, nms_threshold: float = 0.5, class_specific_filter: bool = True):
        super().__init__()
        self.model = model
        self.nms = nms
        self.score_threshold = score_threshold
        self.max_detections = max_detections
        self.nms_threshold = nms_threshold
        self.class_specific_filter = class_specific_filter

    @torch.no_grad()
    def forward(self, images: torch.Tensor):
        from ..ops.anchors import anchors_for_shape
        from ..ops.boxes import bbox_transform_inv, clip_boxes
        from ..ops.filter_detections import filter_detections

        out = self.model(images)
        h, w = images.shape[-2:]
        anchors = anchors_for_shape((h, w), feature_shapes=out["feature_shapes"]).to(images.device)
        boxes = bbox_transform_inv(anchors.unsqueeze(0), out["regression"].float())
        boxes = clip_boxes(boxes, (h, w))
        cls = out["classification"].float()
        scores = torch.sigmoid(cls) if self.model.return_logits else cls
        return filter_detections(boxes, scores, nms=self.nms, score_threshold=self.score_threshold,
                                 max_detections=self.max_detections, nms_threshold=self.nms_threshold,
                                 class_specific_filter=self.class_specific_filter)


def retinanet_bbox(model: RetinaNet, **kw) -> RetinaNetBBox:
    return RetinaNetBBox(model, **kw)


def retinanet_forward_macs(depth: str, h: int, w: int, num_classes: int = 80) -> dict:
    """Forward MACs split by part (244.6 GMAC for R50 @ 800x1333)."""
    from .resnet import resnet_forward_macs

    def c(n, k=1):
        for _ in range(k):
            n = -(-n // 2)
        return n
    bb = resnet_forward_macs(depth, h, w)
    lv = [(c(h, l), c(w, l)) for l in range(3, 8)]
    px = [a * b for a, b in lv]
    f = FEATURE_SIZE
    fpn = px[0] * 512 * f + px[1] * 1024 * f + px[2] * 2048 * f + (px[0] + px[1] + px[2]) * f * f * 9
    fpn += px[3] * 2048 * f * 9 + px[4] * f * f * 9
    tot = sum(px)
    heads = tot * (8 * f * f * 9 + f * 9 * (NUM_ANCHORS * num_classes + NUM_ANCHORS * 4))
    return {"backbone": bb, "fpn": fpn, "heads": heads, "total": bb + fpn + heads}
