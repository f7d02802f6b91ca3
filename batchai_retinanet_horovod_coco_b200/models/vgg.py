"""VGG16 / VGG19 backbones (spec layer, plain PyTorch).

The reference accepts any ``--backbone`` that ``keras_retinanet.models.backbone`` knows and only
warns when it is not a ResNet (``train.py:323-324,390``).  keras-retinanet's VGG
backbone is ``keras.applications.VGG16/19(include_top=False)`` with the pyramid fed from
``block3_pool``, ``block4_pool`` and ``block5_pool`` (strides 8 / 16 / 32, 256 / 512 / 512 channels).

Behaviour pinned here:

* every conv is 3x3, stride 1, ``same`` padding, with bias, followed by ReLU; no normalisation;
* pools are 2x2 / 2 with ``valid`` padding, so feature sizes are **floor** divisions (the pyramid
  shapes come from the tensors, not from the ceil-div guess used for ResNet);
* Keras layer names ``block{b}_conv{i}`` so ``.h5`` weight files map by name;
* all layers are trainable unless ``--freeze-backbone``; preprocessing is caffe mode (BGR mean).
"""

from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

VGG_CFG = {
    "vgg16": (2, 2, 3, 3, 3),
    "vgg19": (2, 2, 4, 4, 4),
}
VGG_WIDTH = (64, 128, 256, 512, 512)


class VGG(nn.Module):
    preprocess_mode = "caffe"

    def __init__(self, name: str = "vgg16"):
        super().__init__()
        if name not in VGG_CFG:
            raise ValueError(f"unknown VGG variant '{name}' (have {sorted(VGG_CFG)})")
        self.name = name
        self.blocks = nn.ModuleList()
        self._names = []
        cin = 3
        for b, (n, width) in enumerate(zip(VGG_CFG[name], VGG_WIDTH), start=1):
            convs = nn.ModuleList()
            for i in range(1, n + 1):
                convs.append(nn.Conv2d(cin, width, 3, padding=1))
                self._names.append(f"block{b}_conv{i}")
                cin = width
            self.blocks.append(convs)
        self.out_channels = [VGG_WIDTH[1], VGG_WIDTH[2], VGG_WIDTH[3], VGG_WIDTH[4]]
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.xavier_uniform_(m.weight)   # Keras glorot_uniform default
                nn.init.zeros_(m.bias)

    def keras_layers(self) -> "OrderedDict[str, nn.Module]":
        d = OrderedDict()
        it = iter(self._names)
        for convs in self.blocks:
            for c in convs:
                d[next(it)] = c
        return d

    @staticmethod
    def feature_hw(h: int, w: int):
        """(h, w) of C3, C4, C5: 2x2/2 ``valid`` pools floor."""
        return [(h >> k, w >> k) for k in (3, 4, 5)]

    def forward(self, x: torch.Tensor):
        feats = []
        for convs in self.blocks:
            for c in convs:
                x = F.relu(c(x))
            x = F.max_pool2d(x, 2, 2)
            feats.append(x)
        return feats[1], feats[2], feats[3], feats[4]
