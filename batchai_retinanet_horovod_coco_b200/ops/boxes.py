"""Box decoding / clipping (spec layer).

Keras-retinanet ``RegressBoxes(mean=0, std=0.2)`` and ``ClipBoxes`` as used by the prediction
graph built at reference ``train.py:95``.
"""

from __future__ import annotations

import torch


def bbox_transform_inv(anchors: torch.Tensor, deltas: torch.Tensor, mean=0.0, std=0.2) -> torch.Tensor:
    """Corner-offset decode: ``x1 = a_x1 + (d*std+mean) * a_w`` etc.  Shapes (...,4)."""
    w = anchors[..., 2] - anchors[..., 0]
    h = anchors[..., 3] - anchors[..., 1]
    d = deltas * std + mean
    return torch.stack([anchors[..., 0] + d[..., 0] * w, anchors[..., 1] + d[..., 1] * h,
                        anchors[..., 2] + d[..., 2] * w, anchors[..., 3] + d[..., 3] * h], dim=-1)


def clip_boxes(boxes: torch.Tensor, image_hw) -> torch.Tensor:
    h, w = image_hw
    return torch.stack([boxes[..., 0].clamp(0, w), boxes[..., 1].clamp(0, h),
                        boxes[..., 2].clamp(0, w), boxes[..., 3].clamp(0, h)], dim=-1)
