"""VGG16 / VGG19 backbone on the native kernels.

A VGG backbone is nothing but 3x3 stride-1 convolutions with bias + ReLU and 2x2 max-pools, which the
tcgen05 implicit-GEMM kernels of the ResNet engine already cover (``csrc/conv_sm100.cu``, ``wgrad_sm100.cu``);
only the pool is new (``b200_maxpool2_fwd`` / ``b200_maxpool2_relu_bwd``).  So unlike MobileNet / DenseNet
(``engine/hybrid.py``) this backbone does not go through autograd: forward, backward, bias gradients and the
flat-buffer optimizer are the same hand-scheduled path as for ResNet, and the whole step is graph-captured.

Details that differ from ResNet:

* the image is preprocessed into an 8-channel NHWC buffer (TMA needs a 16-byte pixel stride); ``block1_conv1``
  runs with Cin = 3 inside a zero-filled 64-wide K tile;
* the weight gradient of that one layer (Cin = 3 cannot be a 64-channel MN-major operand) comes from the small-Cin
  CUDA-core kernel ``b200_wgrad_small``: no library call is left in the step;
* pools floor, so the pyramid shapes come from ``VGG.feature_hw``; C3 / C4 / C5 are ``block{3,4,5}_pool``
  (keras-retinanet's VGG backbone, reachable through ``--backbone vgg16`` at reference ``train.py:390``).
"""

from __future__ import annotations

import torch

from ..ops import native
from .trainer import BF16, CAFFE_MEAN_BGR, Conv, NativeTrainer


class VGGTrainer(NativeTrainer):
    # ------------------------------------------------------------------ structure
    def _build_backbone(self):
        self.stem = None
        self.blocks = []
        bb = self.model.backbone
        layers = list(bb.keras_layers().items())
        self.vgg = []                      # per block: list of Conv
        it = iter(layers)
        for convs in bb.blocks:
            self.vgg.append([Conv(*next(it)) for _ in convs])
        order = []
        for blk in self.vgg[::-1]:         # backward-completion order
            order += blk[::-1]
        return order

    # ------------------------------------------------------------------ per-shape plan
    def _plan_backbone(self, p, B, H, W, buf):
        dev = self.device
        p["img"] = torch.empty(B, H, W, 8, device=dev, dtype=BF16)      # channels 3..7 are zero padding
        acts, pools, gping, gpool = [], [], [], []
        h, w = H, W
        for blk in self.vgg:
            c = blk[0].cout
            acts.append([buf(h, w, c) for _ in blk])
            gping.append((buf(h, w, c), buf(h, w, c)) if not self.freeze_backbone else None)
            h, w = h // 2, w // 2
            pools.append(buf(h, w, c))
            gpool.append(buf(h, w, c) if not self.freeze_backbone else None)
        p["vgg_acts"], p["vgg_pools"], p["vgg_g"], p["vgg_gpool"] = acts, pools, gping, gpool
        if not self.freeze_backbone:
            p["vgg_glat"] = {2: torch.empty_like(pools[2]), 3: torch.empty_like(pools[3])}
        return [tuple(t.shape[1:3]) for t in pools[2:5]]

    # ------------------------------------------------------------------ forward
    def _forward_backbone(self, p, images_u8, img_hw=None):
        native.preprocess(images_u8, p["img"], CAFFE_MEAN_BGR, img_hw)
        x = p["img"]
        for bi, blk in enumerate(self.vgg):
            for ci, c in enumerate(blk):
                self._fprop(c, x, p["vgg_acts"][bi][ci], relu=True)
                x = p["vgg_acts"][bi][ci]
            native.maxpool2_fwd(x, p["vgg_pools"][bi])
            x = p["vgg_pools"][bi]
        self._mark("fwd_backbone")
        return p["vgg_pools"][2], p["vgg_pools"][3], p["vgg_pools"][4]

    # ------------------------------------------------------------------ backward
    def _backward_backbone(self, p):
        f, df = self.fpn, p["dfpn"]
        acts, pools, gpool, glat = p["vgg_acts"], p["vgg_pools"], p["vgg_gpool"], p["vgg_glat"]
        # gradients arriving from the pyramid: C5 <- C5_reduced + P6 (3x3/2), C4 / C3 <- their laterals
        self._dgrad(f["C5_reduced"], df["c5r"], gpool[4])
        self._dgrad(f["P6"], df["p6"], gpool[4], add=gpool[4])
        self._dgrad(f["C4_reduced"], df["m4"], glat[3])
        self._dgrad(f["C3_reduced"], df["m3"], glat[2])
        for bi in range(len(self.vgg) - 1, -1, -1):
            blk = self.vgg[bi]
            g, other = p["vgg_g"][bi]
            # pool backward + the ReLU mask of the block's last conv -> gradient of its pre-activation
            native.maxpool2_relu_bwd(acts[bi][-1], gpool[bi], g)
            for ci in range(len(blk) - 1, -1, -1):
                c = blk[ci]
                x_in = acts[bi][ci - 1] if ci > 0 else (pools[bi - 1] if bi > 0 else p["img"])
                if bi == 0 and ci == 0:
                    self._wgrad_first(c, x_in, g)
                    break
                self._wgrad(c, [x_in], [g])
                if ci > 0:
                    self._dgrad(c, g, other, mask=x_in)              # x_in is the previous conv's ReLU output
                    g, other = other, g
                else:
                    # into the previous block's pool output (no activation in between); C3 / C4 also feed the FPN
                    self._dgrad(c, g, gpool[bi - 1], add=glat.get(bi - 1))
        self._mark("bwd_backbone")

    def _wgrad_first(self, c: Conv, x8, g):
        """block1_conv1: Cin = 3 is below the granularity of the tensor-core wgrad operand (95 % padding): a small
        CUDA-core kernel (``b200_wgrad_small``) accumulates straight into the gradient bucket."""
        if not c.trainable:
            return
        native.wgrad_small(x8, g, c.wview(self.G), 3, c.cout, 3, 3, 1, (1, 1))
        native.colsum(g, c.cout, c.bview(self.G))
