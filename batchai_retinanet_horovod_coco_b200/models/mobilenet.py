"""MobileNet v1 backbones ``mobilenet{128,160,192,224}_{alpha}`` (spec layer, plain PyTorch).

keras-retinanet's MobileNet backbone (reachable from the reference through ``--backbone``,
``train.py:323-324,390``) is ``keras.applications.MobileNet(alpha, include_top=False)``
with the pyramid fed from ``conv_pw_5_relu`` (stride 8, 256a channels), ``conv_pw_11_relu`` (stride
16, 512a) and ``conv_pw_13_relu`` (stride 32, 1024a).  The number in the name is the ImageNet
resolution of the pretrained file only; the network itself is fully convolutional.

Behaviour pinned here:

* ``conv1`` 3x3/2 (no bias) + BN + ReLU6, then 13 blocks of depthwise 3x3 (no bias) + BN + ReLU6 +
  pointwise 1x1 (no bias) + BN + ReLU6; strides 2 at blocks 2, 4, 6, 12; TensorFlow ``same`` padding;
* BatchNorm layers are ordinary **trainable** Keras BN (momentum 0.99, eps 1e-3): batch statistics in
  training, moving statistics at inference, unlike the frozen BN of the ResNet backbones;
* Keras layer names (``conv1``, ``conv1_bn``, ``conv_dw_{i}``, ``conv_dw_{i}_bn``, ``conv_pw_{i}``,
  ``conv_pw_{i}_bn``) so ``.h5`` files map by name; preprocessing is ``tf`` mode (x/127.5 - 1).
"""

from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

from .resnet import tf_same_pad

ALLOWED = [f"mobilenet{r}_{a}" for r in (128, 160, 192, 224) for a in ("1.0", "0.75", "0.5", "0.25")]
# (pointwise filters, stride of the depthwise conv) for blocks 1..13
MOBILENET_BLOCKS = [(64, 1), (128, 2), (128, 1), (256, 2), (256, 1), (512, 2), (512, 1), (512, 1), (512, 1), (512, 1),
                    (512, 1), (1024, 2), (1024, 1)]


class KerasBatchNorm2d(nn.Module):
    """Trainable BatchNorm with Keras names (``gamma``, ``beta``, ``moving_mean``, ``moving_variance``)
    and Keras' momentum convention (moving = moving * m + batch * (1 - m))."""

    def __init__(self, channels: int, eps: float = 1e-3, momentum: float = 0.99):
        super().__init__()
        self.eps = eps
        self.momentum = momentum
        self.gamma = nn.Parameter(torch.ones(channels))
        self.beta = nn.Parameter(torch.zeros(channels))
        self.register_buffer("moving_mean", torch.zeros(channels))
        self.register_buffer("moving_variance", torch.ones(channels))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return F.batch_norm(x, self.moving_mean, self.moving_variance, self.gamma, self.beta, self.training,
                            1.0 - self.momentum, self.eps)


def same_conv(conv: nn.Conv2d, x: torch.Tensor) -> torch.Tensor:
    """Conv with TensorFlow ``same`` padding (asymmetric when the stride is 2)."""
    k, s = conv.kernel_size[0], conv.stride[0]
    ph, pw = tf_same_pad(x.shape[-2], k, s), tf_same_pad(x.shape[-1], k, s)
    x = F.pad(x, (pw[0], pw[1], ph[0], ph[1]))
    return F.conv2d(x, conv.weight, conv.bias, stride=s, groups=conv.groups)


def relu6(x):
    return F.relu6(x)


class MobileNet(nn.Module):
    preprocess_mode = "tf"

    def __init__(self, name: str = "mobilenet224_1.0"):
        super().__init__()
        if name not in ALLOWED:
            raise ValueError(f"Backbone ('{name}') not in allowed backbones ({ALLOWED}).")
        self.name = name
        self.alpha = float(name.split("_")[1])
        a = self.alpha
        c = int(32 * a)
        self.conv1 = nn.Conv2d(3, c, 3, stride=2, bias=False)
        self.conv1_bn = KerasBatchNorm2d(c)
        self.dw = nn.ModuleList()
        self.dw_bn = nn.ModuleList()
        self.pw = nn.ModuleList()
        self.pw_bn = nn.ModuleList()
        for filters, stride in MOBILENET_BLOCKS:
            f = int(filters * a)
            self.dw.append(nn.Conv2d(c, c, 3, stride=stride, groups=c, bias=False))
            self.dw_bn.append(KerasBatchNorm2d(c))
            self.pw.append(nn.Conv2d(c, f, 1, bias=False))
            self.pw_bn.append(KerasBatchNorm2d(f))
            c = f
        self.out_channels = [int(128 * a), int(256 * a), int(512 * a), int(1024 * a)]
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.xavier_uniform_(m.weight)

    def keras_layers(self) -> "OrderedDict[str, nn.Module]":
        d = OrderedDict()
        d["conv1"] = self.conv1
        d["conv1_bn"] = self.conv1_bn
        for i in range(13):
            d[f"conv_dw_{i + 1}"] = self.dw[i]
            d[f"conv_dw_{i + 1}_bn"] = self.dw_bn[i]
            d[f"conv_pw_{i + 1}"] = self.pw[i]
            d[f"conv_pw_{i + 1}_bn"] = self.pw_bn[i]
        return d

    @staticmethod
    def feature_hw(h: int, w: int):
        """(h, w) of C3, C4, C5: every stride-2 conv uses ``same`` padding, so sizes ceil."""
        return [(-(-h // (1 << k)), -(-w // (1 << k))) for k in (3, 4, 5)]

    def forward(self, x: torch.Tensor):
        x = relu6(self.conv1_bn(same_conv(self.conv1, x)))
        outs = {}
        for i in range(13):
            x = relu6(self.dw_bn[i](same_conv(self.dw[i], x)))
            x = relu6(self.pw_bn[i](self.pw[i](x)))
            outs[i + 1] = x
        return outs[3], outs[5], outs[11], outs[13]
