"""MobileNet v1 backbone on the native kernels (no autograd, no cuDNN; the whole step is graph-captured).

keras-retinanet's MobileNet backbone is reachable from the reference through ``--backbone mobilenet224_1.0``
(``train.py:323-324,390``).  Layer types and where they run:

* ``conv1`` (3x3/2, Cin = 3) and the 13 pointwise 1x1 convolutions: the tcgen05 implicit-GEMM kernels of the ResNet
  engine (``csrc/conv_sm100.cu``, ``wgrad_sm100.cu``); ``conv1``'s filter gradient is ``b200_wgrad_small`` (three input
  channels would be 95 % padding as a tensor-core operand);
* the 13 depthwise 3x3 convolutions: ``b200_dwconv3x3`` (forward / input gradient / filter gradient);
* the 27 trainable batch-norm layers with BATCH statistics + ReLU6: ``b200_bn_stats`` / ``b200_bn_act_fwd`` /
  ``b200_bn_act_bwd`` (Keras moving statistics are updated in the forward pass, momentum 0.99; inference and
  ``--freeze-backbone`` use the moving statistics).

Parameters: ``conv1`` and the pointwise convolutions live in the engine's conv table (master layout, bf16 working
copies); depthwise filters and the BN gamma / beta are views into the same flat ``P`` / ``G`` buffers (machinery of
``HybridTrainer``), so the clip, the P2P allreduce and the Adam kernel see one contiguous range.
"""

from __future__ import annotations

import os

import torch

from ..models.resnet import tf_same_pad
from ..ops import native
from .hybrid import HybridTrainer
from .trainer import BF16, Conv, _align4


class _BN:
    """One batch-norm layer: parameters (views of P / G), moving statistics, batch statistics of the current step."""

    def __init__(self, module, device):
        self.m = module
        self.c = module.gamma.numel()
        self.mean = torch.zeros(self.c, device=device)
        self.var = torch.ones(self.c, device=device)


class MobileNetTrainer(HybridTrainer):
    ACT = 2        # ReLU6

    # ------------------------------------------------------------------ structure
    def _build_backbone(self):
        graph = self.use_graph
        super()._build_backbone()
        self.use_graph = graph                      # nothing here goes through autograd
        bb = self.bb
        self.conv1 = Conv("conv1", bb.conv1, stride=2)
        self.pw = [Conv(f"conv_pw_{i + 1}", bb.pw[i]) for i in range(13)]
        self._table_ids = {id(c.module.weight) for c in [self.conv1] + self.pw}
        self.bn1 = _BN(bb.conv1_bn, self.device)
        self.bn_dw = [_BN(m, self.device) for m in bb.dw_bn]
        self.bn_pw = [_BN(m, self.device) for m in bb.pw_bn]
        cmax = max(b.c for b in [self.bn1] + self.bn_dw + self.bn_pw)
        self.bn_sums = torch.zeros(2 * cmax, device=self.device)
        return self.pw[::-1] + [self.conv1]          # backward-completion order

    def _alloc_extra(self, off: int, trainable: bool) -> int:
        for prm in self.bb.parameters():
            if prm.requires_grad != trainable or id(prm) in self._table_ids:
                continue
            self._extra.append((prm, off, prm.numel()))
            off = _align4(off + prm.numel())
        return off

    # ------------------------------------------------------------------ per-shape plan
    def _plan_backbone(self, p, B, H, W, buf):
        dev = self.device
        p["img"] = torch.empty(B, H, W, 8, device=dev, dtype=BF16)      # channels 3..7 are zero padding
        h, w = -(-H // 2), -(-W // 2)
        c = self.conv1.cout
        p["mb_raw0"], p["mb_act0"] = buf(h, w, c), buf(h, w, c)
        blocks = []
        train = not self.freeze_backbone
        for i in range(13):
            s = self.bb.dw[i].stride[0]
            ho, wo = -(-h // s), -(-w // s)
            f = self.pw[i].cout
            blk = {"in_hw": (h, w), "hw": (ho, wo), "stride": s, "pad": (tf_same_pad(h, 3, s)[0], tf_same_pad(w, 3, s)[0]),
                   "raw_dw": buf(ho, wo, c), "act_dw": buf(ho, wo, c), "raw_pw": buf(ho, wo, f), "act_pw": buf(ho, wo, f)}
            if train:
                blk.update({"g_pw": buf(ho, wo, f), "g_dw": buf(ho, wo, c), "g_in": buf(h, w, c)})
            blocks.append(blk)
            h, w, c = ho, wo, f
        p["mb"] = blocks
        if train:
            p["mb_g0"] = buf(*p["mb_raw0"].shape[1:3], self.conv1.cout)
            p["gC"] = [buf(*blocks[4]["hw"], self.pw[4].cout), buf(*blocks[10]["hw"], self.pw[10].cout),
                       buf(*blocks[12]["hw"], self.pw[12].cout)]
        return [blocks[4]["hw"], blocks[10]["hw"], blocks[12]["hw"]]

    # ------------------------------------------------------------------ forward
    def _bn_forward(self, bn: _BN, raw, act, batch_stats: bool):
        m = bn.m
        if batch_stats:
            native.bn_stats(raw, bn.c, self.bn_sums, bn.mean, bn.var, m.moving_mean, m.moving_variance, m.momentum)
            native.bn_act_fwd(raw, bn.mean, bn.var, m.gamma.data, m.beta.data, m.eps, act, bn.c, self.ACT)
        else:
            native.bn_act_fwd(raw, m.moving_mean, m.moving_variance, m.gamma.data, m.beta.data, m.eps, act, bn.c, self.ACT)

    def _forward_backbone(self, p, images_u8, img_hw=None):
        train = not (self._infer or self.freeze_backbone)
        native.preprocess(images_u8, p["img"], (127.5, 127.5, 127.5), img_hw, scale=1.0 / 127.5)       # "tf" mode
        self._fprop(self.conv1, p["img"], p["mb_raw0"])
        self._bn_forward(self.bn1, p["mb_raw0"], p["mb_act0"], train)
        x = p["mb_act0"]
        for i, b in enumerate(p["mb"]):
            native.dwconv3x3(0, x, self.bb.dw[i].weight.data.view(-1, 9), b["raw_dw"], b["in_hw"], b["hw"], b["stride"], b["pad"])
            self._bn_forward(self.bn_dw[i], b["raw_dw"], b["act_dw"], train)
            self._fprop(self.pw[i], b["act_dw"], b["raw_pw"])
            self._bn_forward(self.bn_pw[i], b["raw_pw"], b["act_pw"], train)
            x = b["act_pw"]
        self._mark("fwd_backbone")
        mb = p["mb"]
        return mb[4]["act_pw"], mb[10]["act_pw"], mb[12]["act_pw"]

    # ------------------------------------------------------------------ backward
    def _bn_backward(self, bn: _BN, raw, dy, dx):
        m = bn.m
        native.bn_act_bwd(raw, dy, bn.mean, bn.var, m.gamma.data, m.beta.data, m.eps, self.bn_sums, bn.c, self.ACT, dx,
                          m.gamma.grad if m.gamma.requires_grad else None, m.beta.grad if m.beta.requires_grad else None)

    def _backward_backbone(self, p):
        f, df = self.fpn, p["dfpn"]
        g3, g4, g5 = p["gC"]
        self._dgrad(f["C5_reduced"], df["c5r"], g5)
        self._dgrad(f["P6"], df["p6"], g5, add=g5)
        self._dgrad(f["C4_reduced"], df["m4"], g4)
        self._dgrad(f["C3_reduced"], df["m3"], g3)
        lateral = {4: g3, 10: g4}
        mb = p["mb"]
        g = g5
        for i in range(12, -1, -1):
            b = mb[i]
            x_in = mb[i - 1]["act_pw"] if i > 0 else p["mb_act0"]
            self._bn_backward(self.bn_pw[i], b["raw_pw"], g, b["g_pw"])
            self._wgrad(self.pw[i], [b["act_dw"]], [b["g_pw"]])
            self._dgrad(self.pw[i], b["g_pw"], b["g_dw"])
            self._bn_backward(self.bn_dw[i], b["raw_dw"], b["g_dw"], b["g_dw"])          # in place: same thread reads, then writes
            dwm = self.bb.dw[i].weight
            if dwm.requires_grad:
                native.dwconv3x3(2, x_in, b["g_dw"], dwm.grad.view(-1, 9), b["in_hw"], b["hw"], b["stride"], b["pad"])
            native.dwconv3x3(1, b["g_dw"], dwm.data.view(-1, 9), b["g_in"], b["in_hw"], b["hw"], b["stride"], b["pad"])
            g = b["g_in"]
            if (i - 1) in lateral:          # C3 / C4 also feed the pyramid
                native.add_mask(g, lateral[i - 1], None, g)
        self._bn_backward(self.bn1, p["mb_raw0"], g, p["mb_g0"])
        if self.conv1.trainable:
            H, W = p["img"].shape[1:3]
            native.wgrad_small(p["img"], p["mb_g0"], self.conv1.wview(self.G), 3, self.conv1.cout, 3, 3, 2,
                               (tf_same_pad(H, 3, 2)[0], tf_same_pad(W, 3, 2)[0]))
        self._mark("bwd_backbone")

    def gradients_by_layer(self):
        out = super().gradients_by_layer()
        for c in [self.conv1] + self.pw:          # also under the Keras weight name, like the bound parameters
            if c.name in out:
                out[f"{c.name}/kernel:0"] = out[c.name]
        return out
