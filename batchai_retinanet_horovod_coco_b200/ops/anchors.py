"""Anchor generation and IoU target assignment (spec layer: NumPy-free PyTorch, CPU or GPU).

Behavioural spec: keras-retinanet ``utils/anchors.py`` as reached from the reference through
``anchor_targets_bbox`` (``train.py:51,429``) and ``retinanet_bbox`` (``train.py:95``);
SURVEY.md B7, B11, B12, Appendix C.

* ``AnchorParameters.default``: sizes 32..512, strides 8..128, ratios {0.5,1,2},
  scales {1, 2^(1/3), 2^(2/3)} -> 9 anchors per location, ratio-major order.
* level shapes by ceil-div (``guess_shapes``): ``(s + 2**l - 1) // 2**l``;
* anchors centred at ``(i + 0.5) * stride``;
* assignment: IoU < 0.4 background, >= 0.5 positive, between -> ignore; anchors whose centre
  is outside the un-padded image -> ignore;
* regression targets: corner offsets / anchor (w, h), then (t - 0) / 0.2.

The native GPU path (csrc/targets.cu) implements the same rules; this file is its oracle.
"""

from __future__ import annotations

import functools

import torch

SIZES = [32, 64, 128, 256, 512]
STRIDES = [8, 16, 32, 64, 128]
RATIOS = [0.5, 1.0, 2.0]
SCALES = [2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)]
PYRAMID_LEVELS = [3, 4, 5, 6, 7]


class AnchorParameters:
    def __init__(self, sizes=SIZES, strides=STRIDES, ratios=RATIOS, scales=SCALES):
        self.sizes, self.strides, self.ratios, self.scales = list(sizes), list(strides), list(ratios), list(scales)

    def num_anchors(self) -> int:
        return len(self.ratios) * len(self.scales)


AnchorParameters.default = AnchorParameters()


def generate_anchors(base_size: float = 16, ratios=None, scales=None) -> torch.Tensor:
    """(A,4) reference anchors centred on the origin, ratio-major then scale."""
    ratios = RATIOS if ratios is None else ratios
    scales = SCALES if scales is None else scales
    r = torch.tensor(ratios, dtype=torch.float64).repeat_interleave(len(scales))
    s = torch.tensor(scales, dtype=torch.float64).repeat(len(ratios))
    side = base_size * s
    areas = side * side
    w = torch.sqrt(areas / r)
    h = w * r
    return torch.stack([-w / 2, -h / 2, w / 2, h / 2], dim=1)


def shift(shape, stride: int, anchors: torch.Tensor) -> torch.Tensor:
    """Tile ``anchors`` over an (H, W) grid; row-major pixels, anchor index fastest."""
    h, w = int(shape[0]), int(shape[1])
    sx = (torch.arange(w, dtype=torch.float64) + 0.5) * stride
    sy = (torch.arange(h, dtype=torch.float64) + 0.5) * stride
    yy, xx = torch.meshgrid(sy, sx, indexing="ij")
    shifts = torch.stack([xx.reshape(-1), yy.reshape(-1), xx.reshape(-1), yy.reshape(-1)], dim=1)
    return (anchors.unsqueeze(0) + shifts.unsqueeze(1)).reshape(-1, 4)


def guess_shapes(image_shape, pyramid_levels=PYRAMID_LEVELS):
    h, w = int(image_shape[0]), int(image_shape[1])
    return [((h + 2 ** l - 1) // (2 ** l), (w + 2 ** l - 1) // (2 ** l)) for l in pyramid_levels]


@functools.lru_cache(maxsize=64)
def _anchors_cached(image_shape, feature_shapes, key):
    params = AnchorParameters.default
    out = []
    for i, fs in enumerate(feature_shapes):
        a = generate_anchors(params.sizes[i], params.ratios, params.scales)
        out.append(shift(fs, params.strides[i], a))
    return torch.cat(out, dim=0).to(torch.float32)


def anchors_for_shape(image_shape, feature_shapes=None, anchor_params: AnchorParameters | None = None,
                      shapes_callback=None) -> torch.Tensor:
    """All anchors (sum_l H_l*W_l*9, 4) as float32 x1,y1,x2,y2 for an (H, W) image.

    ``shapes_callback`` mirrors ``make_shapes_callback`` at reference ``train.py:429``.
    """
    image_shape = (int(image_shape[0]), int(image_shape[1]))
    if feature_shapes is None:
        feature_shapes = (shapes_callback or guess_shapes)(image_shape)
    feature_shapes = tuple((int(a), int(b)) for a, b in feature_shapes)
    if anchor_params is None or anchor_params is AnchorParameters.default:
        return _anchors_cached(image_shape, feature_shapes, 0)
    out = []
    for i, fs in enumerate(feature_shapes):
        a = generate_anchors(anchor_params.sizes[i], anchor_params.ratios, anchor_params.scales)
        out.append(shift(fs, anchor_params.strides[i], a))
    return torch.cat(out, dim=0).to(torch.float32)


def compute_overlap(boxes: torch.Tensor, query: torch.Tensor, plus_one: bool = True) -> torch.Tensor:
    """IoU matrix (N,K).  ``plus_one`` reproduces the Cython kernel's +1 pixel convention
    (SURVEY.md B12); ``False`` gives the continuous-coordinate IoU used by NMS."""
    o = 1.0 if plus_one else 0.0
    b = boxes.unsqueeze(1)
    q = query.unsqueeze(0)
    iw = (torch.minimum(b[..., 2], q[..., 2]) - torch.maximum(b[..., 0], q[..., 0]) + o).clamp(min=0)
    ih = (torch.minimum(b[..., 3], q[..., 3]) - torch.maximum(b[..., 1], q[..., 1]) + o).clamp(min=0)
    inter = iw * ih
    area_b = (b[..., 2] - b[..., 0] + o) * (b[..., 3] - b[..., 1] + o)
    area_q = (q[..., 2] - q[..., 0] + o) * (q[..., 3] - q[..., 1] + o)
    ua = (area_b + area_q - inter).clamp(min=1e-8)
    return inter / ua


def bbox_transform(anchors: torch.Tensor, gt: torch.Tensor, mean=0.0, std=0.2) -> torch.Tensor:
    aw = anchors[:, 2] - anchors[:, 0]
    ah = anchors[:, 3] - anchors[:, 1]
    t = torch.stack([(gt[:, 0] - anchors[:, 0]) / aw, (gt[:, 1] - anchors[:, 1]) / ah,
                     (gt[:, 2] - anchors[:, 2]) / aw, (gt[:, 3] - anchors[:, 3]) / ah], dim=1)
    return (t - mean) / std


def assign_targets(anchors: torch.Tensor, annotations: torch.Tensor, image_shape=None,
                   negative_overlap: float = 0.4, positive_overlap: float = 0.5):
    """Compact assignment for one image.

    annotations: (G,5) x1,y1,x2,y2,label.  Returns ``(state (A,) int8 in {-1,0,1},
    matched_gt (A,) int64, max_iou (A,))``.  ``image_shape`` is the un-padded (H, W).
    This is the representation the fused loss kernels consume instead of dense one-hots
    (SURVEY.md G20).
    """
    a = anchors.shape[0]
    dev = anchors.device
    if annotations.numel() == 0:
        state = torch.zeros(a, dtype=torch.int8, device=dev)
        matched = torch.zeros(a, dtype=torch.int64, device=dev)
        max_iou = torch.zeros(a, dtype=torch.float32, device=dev)
    else:
        iou = compute_overlap(anchors.double(), annotations[:, :4].double())
        max_iou, matched = iou.max(dim=1)
        state = torch.full((a,), -1, dtype=torch.int8, device=dev)
        state[max_iou < negative_overlap] = 0
        state[max_iou >= positive_overlap] = 1
        max_iou = max_iou.float()
    if image_shape is not None:
        cx = (anchors[:, 0] + anchors[:, 2]) / 2
        cy = (anchors[:, 1] + anchors[:, 3]) / 2
        outside = (cx >= image_shape[1]) | (cy >= image_shape[0])
        state[outside] = -1
    return state, matched, max_iou


def anchor_targets_bbox(anchors: torch.Tensor, annotations_group, num_classes: int, image_shapes=None,
                        negative_overlap: float = 0.4, positive_overlap: float = 0.5):
    """Dense Keras-format targets for a batch (used by the spec path and by tests).

    Returns ``regression (B,A,5)`` and ``labels (B,A,K+1)``; last channel = anchor state.
    """
    b = len(annotations_group)
    a = anchors.shape[0]
    reg = torch.zeros(b, a, 5)
    lab = torch.zeros(b, a, num_classes + 1)
    for i, ann in enumerate(annotations_group):
        ann = torch.as_tensor(ann, dtype=torch.float32).reshape(-1, 5)
        shape = None if image_shapes is None else image_shapes[i]
        state, matched, _ = assign_targets(anchors, ann, shape, negative_overlap, positive_overlap)
        lab[i, :, -1] = state.float()
        reg[i, :, -1] = state.float()
        if ann.shape[0]:
            pos = state == 1
            g = ann[matched]
            lab[i, pos, g[pos, 4].long()] = 1
            reg[i, :, :4] = bbox_transform(anchors, g[:, :4])
    return reg, lab
