"""Focal and smooth-L1 losses (spec layer).

Behavioural spec: keras-retinanet ``losses.focal(alpha=0.25, gamma=2.0)`` and
``losses.smooth_l1(sigma=3.0)`` bound at reference ``train.py:99-102``.

Two input conventions are provided:

* ``focal`` / ``smooth_l1`` take the dense Keras targets ``(B,A,K+1)`` / ``(B,A,5)`` whose last
  channel is the anchor state (-1 ignore, 0 background, 1 positive) and, for ``focal``,
  **probabilities** clipped to [1e-7, 1-1e-7] exactly as Keras' binary cross-entropy does;
* ``focal_compact`` / ``smooth_l1_compact`` take **logits** plus the compact assignment
  ``(state, class_id)`` and encoded regression targets.  They are the oracle for, and the
  CPU fallback of, the fused CUDA kernels in ``csrc/losses.cu``.

Both normalise by ``max(1, #positive anchors)`` counted over the whole batch.
"""

from __future__ import annotations

import torch
import torch.nn.functional as F


def focal(y_true: torch.Tensor, y_pred: torch.Tensor, alpha: float = 0.25, gamma: float = 2.0) -> torch.Tensor:
    labels = y_true[..., :-1]
    state = y_true[..., -1]
    keep = state != -1
    labels = labels[keep]
    p = y_pred[keep].float()
    alpha_t = torch.where(labels == 1, torch.full_like(p, alpha), torch.full_like(p, 1 - alpha))
    w = torch.where(labels == 1, 1 - p, p)
    w = alpha_t * w ** gamma
    pc = p.clamp(1e-7, 1 - 1e-7)
    bce = -(labels * torch.log(pc) + (1 - labels) * torch.log(1 - pc))
    normalizer = (state == 1).sum().clamp(min=1).float()
    return (w * bce).sum() / normalizer


def smooth_l1(y_true: torch.Tensor, y_pred: torch.Tensor, sigma: float = 3.0) -> torch.Tensor:
    s2 = sigma * sigma
    target = y_true[..., :4]
    state = y_true[..., 4]
    pos = state == 1
    d = (y_pred[pos].float() - target[pos]).abs()
    loss = torch.where(d < 1.0 / s2, 0.5 * s2 * d * d, d - 0.5 / s2)
    normalizer = pos.sum().clamp(min=1).float()
    return loss.sum() / normalizer


def focal_compact(logits: torch.Tensor, state: torch.Tensor, cls: torch.Tensor,
                  alpha: float = 0.25, gamma: float = 2.0, num_pos: torch.Tensor | None = None) -> torch.Tensor:
    """logits (B,A,K); state (B,A) int8; cls (B,A) class id of the matched GT."""
    x = logits.float()
    k = x.shape[-1]
    onehot = (torch.arange(k, device=x.device).view(1, 1, k) == cls.unsqueeze(-1).long()) & (state == 1).unsqueeze(-1)
    t = onehot.float()
    p = torch.sigmoid(x)
    ce = F.binary_cross_entropy_with_logits(x, t, reduction="none")
    p_t = p * t + (1 - p) * (1 - t)
    a_t = alpha * t + (1 - alpha) * (1 - t)
    loss = a_t * (1 - p_t) ** gamma * ce
    loss = loss * (state != -1).unsqueeze(-1).float()
    if num_pos is None:
        num_pos = (state == 1).sum()
    return loss.sum() / num_pos.clamp(min=1).float()


def smooth_l1_compact(pred: torch.Tensor, target: torch.Tensor, state: torch.Tensor, sigma: float = 3.0,
                      num_pos: torch.Tensor | None = None) -> torch.Tensor:
    """pred, target (B,A,4); state (B,A)."""
    s2 = sigma * sigma
    pos = (state == 1).unsqueeze(-1).float()
    d = (pred.float() - target.float()).abs()
    loss = torch.where(d < 1.0 / s2, 0.5 * s2 * d * d, d - 0.5 / s2) * pos
    if num_pos is None:
        num_pos = (state == 1).sum()
    return loss.sum() / num_pos.clamp(min=1).float()
