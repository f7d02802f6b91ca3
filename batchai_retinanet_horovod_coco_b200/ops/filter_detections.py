"""``filter_detections`` (spec layer) -- keras-retinanet ``FilterDetections`` as built into
the prediction graph at reference ``train.py:95``.

Per image, class-specific mode: keep ``score > 0.05``; per-class NMS (IoU 0.5, at most
``max_detections`` per class); union over classes; global top-``max_detections``; pad with -1.
"""

from __future__ import annotations

import torch


def box_iou_matrix(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    area_a = (a[:, 2] - a[:, 0]).clamp(min=0) * (a[:, 3] - a[:, 1]).clamp(min=0)
    area_b = (b[:, 2] - b[:, 0]).clamp(min=0) * (b[:, 3] - b[:, 1]).clamp(min=0)
    lt = torch.maximum(a[:, None, :2], b[None, :, :2])
    rb = torch.minimum(a[:, None, 2:], b[None, :, 2:])
    wh = (rb - lt).clamp(min=0)
    inter = wh[..., 0] * wh[..., 1]
    return inter / (area_a[:, None] + area_b[None, :] - inter).clamp(min=1e-12)


def nms_reference(boxes: torch.Tensor, scores: torch.Tensor, iou_threshold: float, max_output: int | None = None):
    """Greedy NMS in plain PyTorch (oracle for csrc/nms.cu).  Returns kept indices, score-desc."""
    if boxes.numel() == 0:
        return torch.zeros(0, dtype=torch.int64, device=boxes.device)
    order = scores.argsort(descending=True, stable=True)
    b = boxes[order]
    iou = box_iou_matrix(b, b)
    n = b.shape[0]
    suppressed = torch.zeros(n, dtype=torch.bool)
    iou_c = (iou > iou_threshold).cpu()
    keep = []
    for i in range(n):
        if suppressed[i]:
            continue
        keep.append(i)
        if max_output is not None and len(keep) >= max_output:
            break
        suppressed |= iou_c[i]
    return order[torch.tensor(keep, dtype=torch.int64, device=boxes.device)]


def _nms(boxes, scores, thr, max_output):
    if boxes.is_cuda:
        from . import native
        if native.available():
            return native.nms(boxes, scores, thr, max_output)
    return nms_reference(boxes, scores, thr, max_output)


def filter_detections_single(boxes: torch.Tensor, classification: torch.Tensor, nms: bool = True,
                             class_specific_filter: bool = True, score_threshold: float = 0.05,
                             max_detections: int = 300, nms_threshold: float = 0.5):
    """boxes (A,4), classification (A,K) probabilities -> padded (300,4), (300,), (300,)."""
    dev = boxes.device
    if class_specific_filter:
        cand = (classification > score_threshold).nonzero(as_tuple=False)  # (M,2) anchor, class
        a_idx, c_idx = cand[:, 0], cand[:, 1]
        sc = classification[a_idx, c_idx]
        if nms and a_idx.numel():
            # per-class NMS by the class-offset trick: boxes of different classes never overlap
            span = (boxes.max() - boxes.min()).abs() + 1.0
            off = c_idx.to(boxes.dtype).unsqueeze(1) * span
            keep = _nms(boxes[a_idx] + off, sc, nms_threshold, None)
            # TF caps each class at max_detections; emulate by per-class rank
            kc = c_idx[keep]
            if keep.numel() > max_detections:
                ranks = torch.zeros_like(kc)
                for cval in kc.unique():
                    m = kc == cval
                    ranks[m] = torch.arange(int(m.sum()), device=dev)
                keep = keep[ranks < max_detections]
            a_idx, c_idx, sc = a_idx[keep], c_idx[keep], sc[keep]
    else:
        sc_all, c_all = classification.max(dim=1)
        a_idx = (sc_all > score_threshold).nonzero(as_tuple=False)[:, 0]
        sc, c_idx = sc_all[a_idx], c_all[a_idx]
        if nms and a_idx.numel():
            keep = _nms(boxes[a_idx], sc, nms_threshold, max_detections)
            a_idx, c_idx, sc = a_idx[keep], c_idx[keep], sc[keep]
    k = min(max_detections, sc.numel())
    top_sc, top_i = torch.topk(sc, k) if k else (sc[:0], torch.zeros(0, dtype=torch.int64, device=dev))
    out_b = torch.full((max_detections, 4), -1.0, device=dev, dtype=boxes.dtype)
    out_s = torch.full((max_detections,), -1.0, device=dev, dtype=classification.dtype)
    out_l = torch.full((max_detections,), -1, device=dev, dtype=torch.int64)
    out_b[:k] = boxes[a_idx[top_i]]
    out_s[:k] = top_sc
    out_l[:k] = c_idx[top_i]
    return out_b, out_s, out_l


def filter_detections(boxes: torch.Tensor, classification: torch.Tensor, **kw):
    """Batched wrapper: boxes (B,A,4), classification (B,A,K)."""
    if boxes.is_cuda:
        from . import native
        if native.available() and kw.get("nms", True) and kw.get("class_specific_filter", True):
            return native.filter_detections(boxes, classification, kw.get("score_threshold", 0.05),
                                            kw.get("max_detections", 300), kw.get("nms_threshold", 0.5))
    outs = [filter_detections_single(boxes[i], classification[i], **kw) for i in range(boxes.shape[0])]
    return (torch.stack([o[0] for o in outs]), torch.stack([o[1] for o in outs]), torch.stack([o[2] for o in outs]))
