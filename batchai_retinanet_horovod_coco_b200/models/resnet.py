"""ResNet-50/101/152 backbone with frozen BatchNorm (spec layer, plain PyTorch).

Behavioural spec: keras-resnet ``ResNet{50,101,152}(include_top=False, freeze_bn=True)`` as
bound by the reference at ``train.py:43,390``.  Layer names follow the
Keras checkpoint layout (``conv1``, ``bn_conv1``, ``res2a_branch2a``, ``bn2a_branch2a`` ...)
so that ``.h5`` files can be read and written by name.

Structure notes that matter for weight parity:

* all backbone convolutions are bias-free;
* the stride-2 of a down-sampling bottleneck sits on the *first* 1x1 (``branch2a``) and on
  the projection shortcut (``branch1``);
* BatchNorm is frozen: always inference-mode affine ``y = x * scale + shift`` with
  ``scale = gamma / sqrt(var + 1e-5)``; gamma/beta/mean/var get no gradient and no update;
* the stem max-pool uses TensorFlow ``same`` padding (pads bottom/right first when the total
  pad is odd), which differs from PyTorch's symmetric ``padding=1`` for even inputs.
"""

from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

BN_EPS = 1e-5

BLOCKS = {
    "resnet50": [3, 4, 6, 3],
    "resnet101": [3, 4, 23, 3],
    "resnet152": [3, 8, 36, 3],
}


def tf_same_pad(n: int, k: int, s: int) -> tuple[int, int]:
    """TensorFlow ``padding='same'`` amounts (before, after) for one spatial dim."""
    out = -(-n // s)
    total = max((out - 1) * s + k - n, 0)
    before = total // 2
    return before, total - before


class FrozenBatchNorm2d(nn.Module):
    """Inference-only BatchNorm whose four vectors are buffers (never trained).

    Keras names: ``gamma``, ``beta``, ``moving_mean``, ``moving_variance``.
    """

    def __init__(self, channels: int, eps: float = BN_EPS):
        super().__init__()
        self.eps = eps
        self.register_buffer("gamma", torch.ones(channels))
        self.register_buffer("beta", torch.zeros(channels))
        self.register_buffer("moving_mean", torch.zeros(channels))
        self.register_buffer("moving_variance", torch.ones(channels))

    def scale_shift(self) -> tuple[torch.Tensor, torch.Tensor]:
        scale = self.gamma * torch.rsqrt(self.moving_variance + self.eps)
        shift = self.beta - self.moving_mean * scale
        return scale, shift

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        scale, shift = self.scale_shift()
        return x * scale.to(x.dtype).view(1, -1, 1, 1) + shift.to(x.dtype).view(1, -1, 1, 1)


def block_name(stage_idx: int, block_idx: int, num_blocks: int) -> str:
    """Keras-resnet block letter: ``a,b,c..`` normally; ``a,b1,b2..`` for long stages."""
    stage = stage_idx + 2
    if block_idx > 0 and num_blocks > 6:
        return f"{stage}b{block_idx}"
    return f"{stage}{chr(ord('a') + block_idx)}"


class Bottleneck(nn.Module):
    def __init__(self, cin: int, filters: int, stride: int, name: str, project: bool):
        super().__init__()
        self.name = name
        self.stride = stride
        self.project = project
        cout = filters * 4
        self.branch2a = nn.Conv2d(cin, filters, 1, stride=stride, bias=False)
        self.bn2a = FrozenBatchNorm2d(filters)
        self.branch2b = nn.Conv2d(filters, filters, 3, padding=1, bias=False)
        self.bn2b = FrozenBatchNorm2d(filters)
        self.branch2c = nn.Conv2d(filters, cout, 1, bias=False)
        self.bn2c = FrozenBatchNorm2d(cout)
        if project:
            self.branch1 = nn.Conv2d(cin, cout, 1, stride=stride, bias=False)
            self.bn1 = FrozenBatchNorm2d(cout)

    def keras_layers(self) -> "OrderedDict[str, nn.Module]":
        n = self.name
        d: "OrderedDict[str, nn.Module]" = OrderedDict()
        d[f"res{n}_branch2a"] = self.branch2a
        d[f"bn{n}_branch2a"] = self.bn2a
        d[f"res{n}_branch2b"] = self.branch2b
        d[f"bn{n}_branch2b"] = self.bn2b
        d[f"res{n}_branch2c"] = self.branch2c
        d[f"bn{n}_branch2c"] = self.bn2c
        if self.project:
            d[f"res{n}_branch1"] = self.branch1
            d[f"bn{n}_branch1"] = self.bn1
        return d

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = F.relu(self.bn2a(self.branch2a(x)))
        y = F.relu(self.bn2b(self.branch2b(y)))
        y = self.bn2c(self.branch2c(y))
        sc = self.bn1(self.branch1(x)) if self.project else x
        return F.relu(y + sc)


def max_pool_same(x: torch.Tensor, k: int = 3, s: int = 2) -> torch.Tensor:
    ph = tf_same_pad(x.shape[-2], k, s)
    pw = tf_same_pad(x.shape[-1], k, s)
    x = F.pad(x, (pw[0], pw[1], ph[0], ph[1]), value=float("-inf"))
    return F.max_pool2d(x, k, s)


class ResNet(nn.Module):
    """Returns (C2, C3, C4, C5); RetinaNet consumes ``outputs[1:]``."""

    def __init__(self, depth: str = "resnet50"):
        super().__init__()
        if depth not in BLOCKS:
            raise ValueError(f"Backbone ('{depth}') not in allowed backbones ({sorted(BLOCKS)}).")
        self.depth = depth
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.bn_conv1 = FrozenBatchNorm2d(64)
        stages = []
        cin = 64
        for si, nb in enumerate(BLOCKS[depth]):
            filters = 64 * (2 ** si)
            blocks = []
            for bi in range(nb):
                stride = 2 if (bi == 0 and si > 0) else 1
                blocks.append(Bottleneck(cin, filters, stride, block_name(si, bi, nb), project=(bi == 0)))
                cin = filters * 4
            stages.append(nn.Sequential(*blocks))
        self.stages = nn.ModuleList(stages)
        self.out_channels = [256, 512, 1024, 2048]
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def keras_layers(self) -> "OrderedDict[str, nn.Module]":
        d: "OrderedDict[str, nn.Module]" = OrderedDict()
        d["conv1"] = self.conv1
        d["bn_conv1"] = self.bn_conv1
        for stage in self.stages:
            for blk in stage:
                d.update(blk.keras_layers())
        return d

    def forward(self, x: torch.Tensor):
        x = F.relu(self.bn_conv1(self.conv1(x)))
        x = max_pool_same(x)
        outs = []
        for stage in self.stages:
            x = stage(x)
            outs.append(x)
        return tuple(outs)


def resnet_forward_macs(depth: str, h: int, w: int) -> int:
    """Multiply-accumulates of one backbone forward (used by roofline reporting)."""
    def c(n, s=2):
        return -(-n // s)
    macs = 0
    h1, w1 = c(h), c(w)
    macs += h1 * w1 * 64 * 3 * 49
    hh, ww = c(h1), c(w1)
    cin = 64
    for si, nb in enumerate(BLOCKS[depth]):
        f = 64 * 2 ** si
        for bi in range(nb):
            if bi == 0 and si > 0:
                ho, wo = c(hh), c(ww)
            else:
                ho, wo = hh, ww
            px = ho * wo
            macs += px * cin * f + px * f * f * 9 + px * f * f * 4
            if bi == 0:
                macs += px * cin * f * 4
            cin = f * 4
            hh, ww = ho, wo
    return macs
