"""DenseNet121 / 169 / 201 backbones (spec layer, plain PyTorch).

keras-retinanet's DenseNet backbone (reachable from the reference through ``--backbone``,
``train.py:323-324,390``) is ``keras.applications.DenseNet{121,169,201}(include_top=
False)`` with the pyramid fed from the concat outputs of dense blocks 2, 3 and 4
(``conv3_block{n}_concat`` stride 8, ``conv4_block{n}_concat`` stride 16, ``conv5_block{n}_concat``
stride 32): 512 / 1024 / 1024 channels for DenseNet121.

Behaviour pinned here:

* stem: zero-pad 3, ``conv1/conv`` 7x7/2 no bias, ``conv1/bn``, ReLU, zero-pad 1, 3x3/2 max-pool;
* dense layer: BN-ReLU-1x1(128, no bias)-BN-ReLU-3x3(32, no bias), concatenated onto its input;
  transition: BN-ReLU-1x1(half the channels, no bias)-2x2/2 average pool; growth rate 32;
* BatchNorm layers are ordinary **trainable** Keras BN with eps 1.001e-5;
* Keras layer names (``conv2_block1_0_bn``, ``conv2_block1_1_conv``, ..., ``pool2_bn``, ``pool2_conv``)
  so ``.h5`` files map by name; preprocessing is ``tf`` mode (x/127.5 - 1).
"""

from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

from .mobilenet import KerasBatchNorm2d

DENSENET_BLOCKS = {
    "densenet121": (6, 12, 24, 16),
    "densenet169": (6, 12, 32, 32),
    "densenet201": (6, 12, 48, 32),
}
GROWTH = 32
DN_EPS = 1.001e-5


class DenseLayer(nn.Module):
    def __init__(self, cin: int):
        super().__init__()
        self.bn0 = KerasBatchNorm2d(cin, eps=DN_EPS)
        self.conv1 = nn.Conv2d(cin, 4 * GROWTH, 1, bias=False)
        self.bn1 = KerasBatchNorm2d(4 * GROWTH, eps=DN_EPS)
        self.conv2 = nn.Conv2d(4 * GROWTH, GROWTH, 3, padding=1, bias=False)

    def forward(self, x):
        y = self.conv1(F.relu(self.bn0(x)))
        y = self.conv2(F.relu(self.bn1(y)))
        return torch.cat([x, y], dim=1)


class Transition(nn.Module):
    def __init__(self, cin: int):
        super().__init__()
        self.bn = KerasBatchNorm2d(cin, eps=DN_EPS)
        self.conv = nn.Conv2d(cin, cin // 2, 1, bias=False)

    def forward(self, x):
        return F.avg_pool2d(self.conv(F.relu(self.bn(x))), 2, 2)


class DenseNet(nn.Module):
    preprocess_mode = "tf"

    def __init__(self, name: str = "densenet121"):
        super().__init__()
        if name not in DENSENET_BLOCKS:
            raise ValueError(f"Backbone ('{name}') not in allowed backbones ({sorted(DENSENET_BLOCKS)}).")
        self.name = name
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = KerasBatchNorm2d(64, eps=DN_EPS)
        c = 64
        self.dense = nn.ModuleList()
        self.trans = nn.ModuleList()
        outs = []
        for bi, n in enumerate(DENSENET_BLOCKS[name]):
            layers = nn.ModuleList()
            for _ in range(n):
                layers.append(DenseLayer(c))
                c += GROWTH
            self.dense.append(layers)
            outs.append(c)
            if bi < 3:
                self.trans.append(Transition(c))
                c //= 2
        self.out_channels = outs
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.xavier_uniform_(m.weight)

    def keras_layers(self) -> "OrderedDict[str, nn.Module]":
        d = OrderedDict()
        d["conv1/conv"] = self.conv1
        d["conv1/bn"] = self.bn1
        for bi, layers in enumerate(self.dense):
            for li, l in enumerate(layers):
                p = f"conv{bi + 2}_block{li + 1}"
                d[f"{p}_0_bn"] = l.bn0
                d[f"{p}_1_conv"] = l.conv1
                d[f"{p}_1_bn"] = l.bn1
                d[f"{p}_2_conv"] = l.conv2
            if bi < 3:
                d[f"pool{bi + 2}_bn"] = self.trans[bi].bn
                d[f"pool{bi + 2}_conv"] = self.trans[bi].conv
        return d

    @staticmethod
    def feature_hw(h: int, w: int):
        """(h, w) of C3, C4, C5: padded stem conv and pool ceil, the 2x2 average pools floor."""
        h, w = -(-h // 2), -(-w // 2)
        h, w = -(-h // 2), -(-w // 2)
        return [(h >> k, w >> k) for k in (1, 2, 3)]

    def forward(self, x: torch.Tensor):
        x = F.relu(self.bn1(self.conv1(x)))
        x = F.max_pool2d(F.pad(x, (1, 1, 1, 1)), 3, 2)
        feats = []
        for bi, layers in enumerate(self.dense):
            for l in layers:
                x = l(x)
            feats.append(x)
            if bi < 3:
                x = self.trans[bi](x)
        return tuple(feats)
