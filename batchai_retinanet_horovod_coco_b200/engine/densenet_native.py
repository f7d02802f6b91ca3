"""DenseNet121 / 169 / 201 backbone on the native kernels (no autograd, no cuDNN; the whole step is graph-captured).

keras-retinanet's DenseNet backbone is reachable from the reference through ``--backbone densenet121``
(``train.py:323-324,390``).  How its layer types map onto kernels:

* concatenation is pointer arithmetic: a dense block owns ONE buffer ``[B, h, w, C_final]``; layer ``l`` reads the channel
  prefix ``[:c_l]`` (channel stride ``C_final``) and its 3x3 convolution stores its 32 new channels straight into
  ``[c_l : c_l + 32]`` through the conv kernel's output tensor map.  The gradient of the block is one buffer of the same
  shape in which every layer ACCUMULATES (``b200_bn_act_bwd(accumulate=1)``) the gradient of the prefix it read;
* BN (batch statistics, trainable) + ReLU in front of every convolution: ``b200_bn_stats`` / ``b200_bn_act_fwd`` /
  ``b200_bn_act_bwd`` on channel-slice views;
* 1x1 (c_l -> 128, C -> C/2) and 3x3 (128 -> 32) convolutions: the tcgen05 kernels (``conv_sm100.cu``, ``wgrad_sm100.cu``;
  c_l = 64 + 32 l does not fill its last 64-channel chunk, which TMA zero-fills);
* transition 2x2 average pool: ``b200_avgpool2`` writing into the next block's buffer prefix;
* stem: the ResNet stem kernels (7x7/2, Cin = 3) without the fused ReLU, then BN + ReLU, then the 3x3/2 max pool.
"""

from __future__ import annotations

import torch

from ..ops import native
from .hybrid import HybridTrainer
from .mobilenet_native import _BN
from .trainer import BF16, Conv, _align4

GROWTH = 32


class DenseNetTrainer(HybridTrainer):
    # ------------------------------------------------------------------ structure
    def _build_backbone(self):
        graph = self.use_graph
        super()._build_backbone()
        self.use_graph = graph
        bb = self.bb
        dev = self.device
        self.dn_stem = Conv("conv1/conv", bb.conv1, stride=2)
        self.dn_bn1 = _BN(bb.bn1, dev)
        self.dn_blocks, self.dn_trans = [], []
        names = iter(bb.keras_layers().items())
        convs = []
        for bi, layers in enumerate(bb.dense):
            blk = []
            for li, l in enumerate(layers):
                p = f"conv{bi + 2}_block{li + 1}"
                d = {"bn0": _BN(l.bn0, dev), "c1": Conv(f"{p}_1_conv", l.conv1), "bn1": _BN(l.bn1, dev),
                     "c2": Conv(f"{p}_2_conv", l.conv2), "cin": l.conv1.weight.shape[1]}
                blk.append(d)
                convs += [d["c1"], d["c2"]]
            self.dn_blocks.append(blk)
            if bi < 3:
                t = bb.trans[bi]
                d = {"bn": _BN(t.bn, dev), "conv": Conv(f"pool{bi + 2}_conv", t.conv)}
                self.dn_trans.append(d)
                convs.append(d["conv"])
        self._table_ids = {id(c.module.weight) for c in convs + [self.dn_stem]}
        cmax = max(b.c for b in self._all_bn())
        self.bn_sums = torch.zeros(2 * cmax, device=dev)
        self.stem = None
        return convs[::-1] + [self.dn_stem]           # backward-completion order

    def _all_bn(self):
        out = [self.dn_bn1]
        for blk in self.dn_blocks:
            for d in blk:
                out += [d["bn0"], d["bn1"]]
        return out + [t["bn"] for t in self.dn_trans]

    def _alloc_extra(self, off: int, trainable: bool) -> int:
        for prm in self.bb.parameters():
            if prm.requires_grad != trainable or id(prm) in self._table_ids:
                continue
            self._extra.append((prm, off, prm.numel()))
            off = _align4(off + prm.numel())
        return off

    def _refresh_weights(self):
        super()._refresh_weights()
        s = self.dn_stem        # packed [64][256] filter of the 7x7 stem kernel (no BN scale to fold: the BN is trainable)
        if not hasattr(self, "dn_stem_wp"):
            self.dn_stem_wp = torch.zeros(64, 256, device=self.device, dtype=BF16)
        self.dn_stem_wp.copy_(native.pack_stem_weight(s.wview(self.P)))

    # ------------------------------------------------------------------ per-shape plan
    def _plan_backbone(self, p, B, H, W, buf):
        dev = self.device
        train = not self.freeze_backbone
        p["img"] = torch.empty(B, H, W, 4, device=dev, dtype=BF16)
        h1, w1 = -(-H // 2), -(-W // 2)
        h2, w2 = (h1 + 2 - 3) // 2 + 1, (w1 + 2 - 3) // 2 + 1
        p["dn_raw0"], p["dn_act0"] = buf(h1, w1, 64), buf(h1, w1, 64)
        p["dn_pool"], p["dn_pool_idx"] = buf(h2, w2, 64), torch.empty(B, h2, w2, 64, device=dev, dtype=torch.uint8)
        if train:
            p["dn_gpool"], p["dn_gact0"], p["dn_graw0"] = buf(h2, w2, 64), buf(h1, w1, 64), buf(h1, w1, 64)
        blocks = []
        h, w, c = h2, w2, 64
        for bi, blk in enumerate(self.dn_blocks):
            cf = c + GROWTH * len(blk)
            d = {"hw": (h, w), "cat": buf(h, w, cf), "layers": []}
            if train:
                d["gcat"] = buf(h, w, cf)
                d["g1"], d["g0"] = buf(h, w, 4 * GROWTH), buf(h, w, cf)       # scratch shared by the block's layers
            for li, l in enumerate(blk):
                cl = c + GROWTH * li
                d["layers"].append({"t0": buf(h, w, cl), "y1": buf(h, w, 4 * GROWTH), "t1": buf(h, w, 4 * GROWTH)})
            if bi < 3:
                d["tt"], d["traw"] = buf(h, w, cf), buf(h, w, cf // 2)
                if train:
                    d["gtraw"] = buf(h, w, cf // 2)
                h, w, c = h // 2, w // 2, cf // 2
            blocks.append(d)
        p["dn"] = blocks
        return [blocks[1]["hw"], blocks[2]["hw"], blocks[3]["hw"]]

    # ------------------------------------------------------------------ helpers
    def _bn_fwd(self, bn: _BN, x, y, act, batch_stats):
        m = bn.m
        if batch_stats:
            native.bn_stats(x, bn.c, self.bn_sums, bn.mean, bn.var, m.moving_mean, m.moving_variance, m.momentum)
            native.bn_act_fwd(x, bn.mean, bn.var, m.gamma.data, m.beta.data, m.eps, y, bn.c, act)
        else:
            native.bn_act_fwd(x, m.moving_mean, m.moving_variance, m.gamma.data, m.beta.data, m.eps, y, bn.c, act)

    def _bn_bwd(self, bn: _BN, x, dy, dx, act, accumulate=False):
        m = bn.m
        native.bn_act_bwd(x, dy, bn.mean, bn.var, m.gamma.data, m.beta.data, m.eps, self.bn_sums, bn.c, act, dx,
                          m.gamma.grad if m.gamma.requires_grad else None, m.beta.grad if m.beta.requires_grad else None,
                          accumulate=accumulate)

    @staticmethod
    def _ld(t):
        return t.stride(-2)

    def _conv(self, c: Conv, x, out):
        """fprop with explicit channel strides (x / out may be channel slices of a concatenation buffer)."""
        native.conv_raw(0, [(x, out, None, None)], c.wf, c.cin, self._ld(x), c.cout, self._ld(out), c.kh, c.kw, 1,
                        c.kh // 2, c.kw // 2)

    def _conv_dgrad(self, c: Conv, dy, out):
        native.conv_raw(1, [(dy, out, None, None)], c.wd, c.cout, self._ld(dy), c.cin, self._ld(out), c.kh, c.kw, 1,
                        c.kh // 2, c.kw // 2)

    def _conv_wgrad(self, c: Conv, x, dy):
        if c.trainable:
            native.wgrad_raw([(x, dy)], c.wview(self.G), c.cin, self._ld(x), c.cout, self._ld(dy), c.kh, c.kw, 1, c.kh // 2,
                             c.kw // 2)

    # ------------------------------------------------------------------ forward
    def _forward_backbone(self, p, images_u8, img_hw=None):
        train = not (self._infer or self.freeze_backbone)
        native.preprocess(images_u8, p["img"], (127.5, 127.5, 127.5), img_hw, scale=1.0 / 127.5)       # "tf" mode
        if not hasattr(self, "_zeros64"):
            self._zeros64, self._ones64 = torch.zeros(64, device=self.device), torch.ones(64, device=self.device)
        native.stem_conv(p["img"], self.dn_stem_wp, self._zeros64, p["dn_raw0"], relu=False)
        self._bn_fwd(self.dn_bn1, p["dn_raw0"], p["dn_act0"], 1, train)
        native.bn_relu_maxpool_fwd(p["dn_act0"], self._ones64, self._zeros64, p["dn_pool"], p["dn_pool_idx"], 1, 1)
        blocks = p["dn"]
        blocks[0]["cat"][..., :64].copy_(p["dn_pool"])
        for bi, (blk, d) in enumerate(zip(self.dn_blocks, blocks)):
            cat = d["cat"]
            for l, bufs in zip(blk, d["layers"]):
                cl = l["cin"]
                self._bn_fwd(l["bn0"], cat[..., :cl], bufs["t0"], 1, train)
                self._conv(l["c1"], bufs["t0"], bufs["y1"])
                self._bn_fwd(l["bn1"], bufs["y1"], bufs["t1"], 1, train)
                self._conv(l["c2"], bufs["t1"], cat[..., cl:cl + GROWTH])          # "concatenate"
            if bi < 3:
                t = self.dn_trans[bi]
                self._bn_fwd(t["bn"], cat, d["tt"], 1, train)
                self._conv(t["conv"], d["tt"], d["traw"])
                nxt = blocks[bi + 1]["cat"]
                native.avgpool2(d["traw"], nxt[..., :d["traw"].shape[-1]])
        self._mark("fwd_backbone")
        return blocks[1]["cat"], blocks[2]["cat"], blocks[3]["cat"]

    # ------------------------------------------------------------------ backward
    def _backward_backbone(self, p):
        f, df = self.fpn, p["dfpn"]
        blocks = p["dn"]
        # the pyramid's gradients land in the concat-gradient buffers of blocks 4, 3, 2
        g5, g4, g3 = blocks[3]["gcat"], blocks[2]["gcat"], blocks[1]["gcat"]
        self._dgrad(f["C5_reduced"], df["c5r"], g5)
        self._dgrad(f["P6"], df["p6"], g5, add=g5)
        self._dgrad(f["C4_reduced"], df["m4"], g4)
        self._dgrad(f["C3_reduced"], df["m3"], g3)
        for bi in range(3, -1, -1):
            blk, d = self.dn_blocks[bi], blocks[bi]
            cat, gcat = d["cat"], d["gcat"]
            if bi < 3:
                # transition: avg-pool, 1x1 conv, BN + ReLU; blocks 2 and 3 already hold the pyramid's gradient
                t = self.dn_trans[bi]
                nxt = blocks[bi + 1]["gcat"]
                native.avgpool2(nxt[..., :d["traw"].shape[-1]], d["gtraw"], backward=True)
                self._conv_wgrad(t["conv"], d["tt"], d["gtraw"])
                self._conv_dgrad(t["conv"], d["gtraw"], d["g0"])
                self._bn_bwd(t["bn"], cat, d["g0"], gcat, 1, accumulate=bi in (1, 2))
            for l, bufs in zip(blk[::-1], d["layers"][::-1]):
                cl = l["cin"]
                g2 = gcat[..., cl:cl + GROWTH]
                self._conv_wgrad(l["c2"], bufs["t1"], g2)
                self._conv_dgrad(l["c2"], g2, d["g1"])
                self._bn_bwd(l["bn1"], bufs["y1"], d["g1"], d["g1"], 1)
                self._conv_wgrad(l["c1"], bufs["t0"], d["g1"])
                g0 = d["g0"][..., :cl]
                self._conv_dgrad(l["c1"], d["g1"], g0)
                self._bn_bwd(l["bn0"], cat[..., :cl], g0, gcat[..., :cl], 1, accumulate=True)
        p["dn_gpool"].copy_(blocks[0]["gcat"][..., :64])
        native.bn_relu_maxpool_bwd(p["dn_pool_idx"], self._ones64, p["dn_gpool"], p["dn_gact0"], 1, 1)
        self._bn_bwd(self.dn_bn1, p["dn_raw0"], p["dn_gact0"], p["dn_graw0"], 1)
        if self.dn_stem.trainable:
            native.stem_wgrad(p["img"], p["dn_graw0"], self._ones64, self.dn_stem.wview(self.G))
        self._mark("bwd_backbone")

    def gradients_by_layer(self):
        out = super().gradients_by_layer()
        for c in self.convs:
            if c.name in out and "/" not in c.name:
                out[f"{c.name}/kernel:0"] = out[c.name]
        if "conv1/conv" in out:
            out["conv1/conv/kernel:0"] = out["conv1/conv"]
        return out
