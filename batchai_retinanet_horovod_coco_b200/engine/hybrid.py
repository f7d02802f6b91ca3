"""Base class of the non-ResNet engines and fallback for a backbone without a native engine.

Since round 2 every backbone family keras-retinanet offers has its own engine on native kernels (``vgg_native.py``,
``mobilenet_native.py``, ``densenet_native.py``); they derive from this class for the flat-buffer plumbing (backbone
parameters and gradients as views of ``P`` / ``G``, batch-norm buffers in the broadcast state, checkpoint names).  Used
directly, the class runs an arbitrary backbone module through autograd as described below.

The reference only warns when ``--backbone`` is not a ResNet (``train.py:323-324``) and then trains
it through the same Keras graph.  Here the part that is identical for every backbone, and that
carries most of the FLOPs -- FPN, both heads over five levels, target assignment, focal / smooth-L1,
clip + allreduce + Adam -- stays on the hand-written sm_100a kernels of ``NativeTrainer``; only the
backbone itself (depthwise convs, trainable batch-norm, dense concatenations: layers the ResNet
flagship does not have) runs through PyTorch autograd in bf16 channels-last.

The seam is three NHWC bf16 tensors (C3, C4, C5) going up and their three gradients coming back:

* backbone parameters are fp32 **views into the flat master buffer** ``P`` and their ``.grad`` are
  views into the flat gradient bucket ``G``, so the norm clip, the P2P allreduce and the Adam kernel
  see one contiguous range exactly as for ResNet (no per-tensor optimizer, no second allreduce);
* batch-norm moving statistics are module buffers; they are broadcast with the rest of the state
  (Horovod ``BroadcastGlobalVariablesCallback(0)`` covers *all* global variables, ``train.py:111``).
"""

from __future__ import annotations

import torch

from .trainer import BF16, NativeTrainer, _align4


class HybridTrainer(NativeTrainer):
    # ------------------------------------------------------------------ structure
    def _build_backbone(self):
        self.stem = None
        self.blocks = []
        self.bb = self.model.backbone.to(self.device)
        self._infer = False
        self.use_graph = False    # the autograd backbone is not captured
        self._extra = []          # (param, offset, numel) in the flat buffers
        self._extra_bound = False
        return []

    def _alloc_extra(self, off: int, trainable: bool) -> int:
        for prm in self.bb.parameters():
            if prm.requires_grad != trainable:
                continue
            self._extra.append((prm, off, prm.numel()))
            off = _align4(off + prm.numel())
        return off

    @torch.no_grad()
    def _bind_extra(self):
        """Make every backbone parameter (and its gradient) a view of the flat buffers."""
        for prm, off, n in self._extra:
            view = self.P[off:off + n].view(prm.shape)
            if not self._extra_bound:
                view.copy_(prm.data.to(self.device, torch.float32))
            prm.data = view
            if prm.requires_grad:
                prm.grad = self.G[off:off + n].view(prm.shape)
        self._extra_bound = True

    # ------------------------------------------------------------------ plan / forward / backward
    def _plan_backbone(self, p, B, H, W, buf):
        hw = self.bb.feature_hw(H, W)
        c3, c4, c5 = self.bb.out_channels[1:]
        p["gC"] = [buf(*hw[0], c3), buf(*hw[1], c4), buf(*hw[2], c5)]
        return hw

    def _forward_backbone(self, p, images_u8, img_hw=None):
        x = images_u8.permute(0, 3, 1, 2).float()          # NCHW view of NHWC memory == channels_last
        if getattr(self.bb, "preprocess_mode", "caffe") == "tf":
            x = x / 127.5 - 1.0
        else:
            x = x - self.mean_t.view(1, 3, 1, 1)
        train = not (self._infer or self.freeze_backbone)
        self.bb.train(train)
        with torch.set_grad_enabled(train), torch.autocast("cuda", dtype=BF16):
            feats = self.bb(x.contiguous(memory_format=torch.channels_last))[1:]
            outs = tuple(f.permute(0, 2, 3, 1).to(BF16).contiguous() for f in feats)   # NHWC, still in the graph
        self._mark("fwd_backbone")
        return outs

    def _backward_backbone(self, p):
        f, df = self.fpn, p["dfpn"]
        g3, g4, g5 = p["gC"]
        self._dgrad(f["C5_reduced"], df["c5r"], g5)
        self._dgrad(f["P6"], df["p6"], g5, add=g5)
        self._dgrad(f["C4_reduced"], df["m4"], g4)
        self._dgrad(f["C3_reduced"], df["m3"], g3)
        torch.autograd.backward(list(p["C"]), [g3, g4, g5])   # accumulates in place into the views of G
        p["C"] = None
        self._mark("bwd_backbone")

    # ------------------------------------------------------------------ public API differences
    def train_step(self, images_u8, gt, gt_count, img_hw):
        self._infer = False
        return super().train_step(images_u8, gt, gt_count, img_hw)

    def forward_loss(self, images_u8, gt, gt_count, img_hw):
        self._infer = False
        return super().forward_loss(images_u8, gt, gt_count, img_hw)

    def predict(self, images_u8, **kw):
        self._infer = True
        try:
            return super().predict(images_u8, **kw)
        finally:
            self._infer = False

    def broadcast_state(self, root: int = 0):
        super().broadcast_state(root)
        if self.comm is None or self.comm.world == 1:
            return
        for b in self.bb.buffers():           # batch-norm moving statistics
            if b.is_floating_point():
                self.comm.broadcast_(b, root)

    def _extra_names(self):
        from ..utils import keras_h5
        return {id(t): n for n, t in keras_h5.trainable_weight_list(self.model)}

    def gradients_by_layer(self):
        out = super().gradients_by_layer()
        names = self._extra_names()
        for prm, off, n in self._extra:
            if prm.requires_grad:
                out[names.get(id(prm), f"extra@{off}")] = (self.G[off:off + n].view(prm.shape).clone(), None)
        return out

    def optimizer_slots(self) -> dict:
        out = super().optimizer_slots()
        names = self._extra_names()
        for prm, off, n in self._extra:
            if prm.requires_grad and id(prm) in names:
                out[names[id(prm)]] = (self.M[off:off + n].view(prm.shape).cpu().clone(),
                                       self.V[off:off + n].view(prm.shape).cpu().clone())
        return out

    @torch.no_grad()
    def load_optimizer_state(self, state: dict):
        super().load_optimizer_state(state)
        slots = state.get("slots", {})
        names = self._extra_names()
        for prm, off, n in self._extra:
            s = slots.get(names.get(id(prm)))
            if s is not None and prm.requires_grad:
                self.M[off:off + n].view(prm.shape).copy_(s[0].to(self.device, torch.float32))
                self.V[off:off + n].view(prm.shape).copy_(s[1].to(self.device, torch.float32))
