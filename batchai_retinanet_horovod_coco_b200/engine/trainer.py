"""Native B200 training/inference engine for RetinaNet (the hot path of this framework).

A static, hand-scheduled executor instead of autograd: the RetinaNet graph is small and fixed,
so forward, losses, backward and the optimizer are an explicit list of launches of the
sm_100a kernels in ``csrc/`` over preallocated NHWC bf16 activations.

Memory plan (all persistent, sized once per (batch, H, W)):

* ``P``  fp32 flat master parameters.  Conv weights are stored ``[KH*KW, Cout, Cin]`` (the layout
  wgrad writes and the working-weight transform reads); biases follow their weights.
* ``G``  fp32 flat gradients, same offsets as ``P``.  wgrad accumulates straight into it
  (``red.global.add.v4.f32``), shared heads sum their five levels there, and it IS the allreduce
  bucket (peer-mapped with CUDA IPC when distributed): no fusion-buffer copies.
* ``M``, ``V`` fp32 Adam slots; ``Wf``/``Wd`` bf16 working weights (BN scale folded) rebuilt by one
  table-driven kernel after each optimizer step.

Step order (reference ``train.py:98-105`` + Horovod ``DistributedOptimizer`` semantics,
SURVEY.md 3.3): forward -> focal + smooth-L1 (fused fwd+bwd) -> backward -> sum g^2 ->
local global-norm clip factor -> allreduce with the per-rank clip factor and 1/N folded into
the peer reads -> Keras-Adam -> working weights.

The PyTorch ``RetinaNet`` module (``models/retinanet.py``) stays the source of truth for
checkpoints: ``load_from_model`` / ``store_to_model`` convert between the two parameter layouts.
"""

from __future__ import annotations

import math
import os
import warnings

import torch

from ..models.resnet import tf_same_pad
from ..models.retinanet import RetinaNet
from ..ops import native
from ..ops.conv import round8

CAFFE_MEAN_BGR = (103.939, 116.779, 123.68)
BF16 = torch.bfloat16


def _align4(n: int) -> int:
    return (n + 3) // 4 * 4


class Conv:
    """One convolution of the model: geometry, flat-buffer offsets, working weights."""

    def __init__(self, name, module, bn=None, stride=1, same=False):
        w = module.weight
        self.name = name
        self.module = module
        self.bn = bn
        self.cout, self.cin, self.kh, self.kw = w.shape
        self.stride = stride
        self.same = same  # TensorFlow 'same' padding (stride-2 3x3)
        self.has_bias = module.bias is not None
        self.taps = self.kh * self.kw
        self.w_off = self.b_off = -1
        self.trainable = w.requires_grad
        self.ldo = round8(self.cout)

    def pads(self, h, w):
        """(pad_before_h, pad_before_w, out_h, out_w)."""
        if self.kh == 1:
            return 0, 0, -(-h // self.stride), -(-w // self.stride)
        if self.stride == 1:
            return self.kh // 2, self.kw // 2, h, w
        ph, pw = tf_same_pad(h, self.kh, self.stride), tf_same_pad(w, self.kw, self.stride)
        return ph[0], pw[0], -(-h // self.stride), -(-w // self.stride)

    # views into the flat buffers ------------------------------------------------------------
    def wview(self, flat):
        return flat[self.w_off:self.w_off + self.taps * self.cout * self.cin].view(self.taps, self.cout, self.cin)

    def bview(self, flat):
        return flat[self.b_off:self.b_off + self.cout]


class StepHandle:
    """Result of ``train_step_async``.  ``loss`` is a device tensor (stream-ordered, always valid);
    ``loss_cpu()`` returns the same triple from pinned host memory after waiting only for the forward half of the
    step -- call it before issuing the next step (a later call falls back to the device tensor)."""

    def __init__(self, trainer):
        self._t = trainer
        self._step = trainer._steps_issued
        self.loss = trainer.loss_out.clone()

    def loss_cpu(self) -> torch.Tensor:
        t = self._t
        if t._steps_issued != self._step:
            return self.loss.cpu()
        t._ev_front.synchronize()
        return t.loss_host.clone()


class NativeTrainer:
    def __init__(self, backbone: str = "resnet50", num_classes: int = 80, device="cuda", batch: int = 8,
                 image_hw=(800, 1333), lr: float = 1e-5, clipnorm: float = 0.001, distributed: bool = False,
                 model: RetinaNet | None = None, freeze_backbone: bool = False, comm=None, shape_bucket: int = 1,
                 max_plans: int = 4, use_graph: bool | None = None):
        native.require()
        self.device = torch.device(device)
        self.num_classes = num_classes
        self.lr = lr
        self.clipnorm = clipnorm
        self.beta1, self.beta2, self.eps = 0.9, 0.999, 1e-7
        self.iterations = 0
        self.freeze_backbone = freeze_backbone
        # Real batches are padded to their own max HxW (reference generator), so shapes change every step:
        # sizes are rounded up to `shape_bucket` and only the `max_plans` most recent buffer plans are kept.
        self.shape_bucket = max(1, int(shape_bucket))
        self.max_plans = max(1, int(max_plans))
        # Whole-step CUDA graph per plan: at small per-GPU batches (the reference's default --batch-size is 1) the
        # ~270 launches of a step cost more host time than GPU time.  B200_GRAPH=0 disables it.
        if use_graph is None:
            use_graph = os.environ.get("B200_GRAPH", "1") != "0"
        self.use_graph = bool(use_graph)
        self.graph_warmup = 2        # eager steps on a plan before it is captured
        self.graph_gt_cap = 128      # ground-truth boxes per image the captured step can take
        self.model = model if model is not None else RetinaNet(num_classes, backbone)
        self.num_classes = self.model.num_classes      # a model that is passed in decides, not the default argument
        if freeze_backbone:
            self.model.freeze_backbone()
        self.backbone_name = self.model.backbone_name
        self.comm = comm
        self.distributed = distributed
        self._build_layers()
        self._alloc_params()
        self.load_from_model(self.model)
        self._plans = {}
        self._prof = None  # list of (name, cuda event) when profiling
        self._comm_prof = None   # [(lo, hi, start event, end event)] of the scatter pieces when tracing
        self._nvtx = os.environ.get("B200_NVTX", "0") == "1"
        # the store-warp column sums merge per-CTA partials with atomics: the deterministic mode uses the ordered
        # column-sum kernel instead
        self._fused_bias = not native.deterministic()
        self.mean_t = torch.tensor(CAFFE_MEAN_BGR, device=self.device)
        if distributed and comm is None:
            from ..parallel.p2p import P2PComm
            self.comm = P2PComm.from_torch_distributed(self.device)
        if self.comm is not None and self.comm.world > 1:
            self._attach_comm()

    # ------------------------------------------------------------------ structure
    def _build_layers(self):
        self.convs = self._build_top() + self._build_backbone()

    def _build_backbone(self):
        """ResNet convs in backward-completion order (stages 5..2, then the stem)."""
        bb = self.model.backbone
        self.stem = Conv("conv1", bb.conv1, bb.bn_conv1, stride=2)
        self.blocks = []  # list of stages, each a list of dicts
        for stage in bb.stages:
            st = []
            for blk in stage:
                d = {"a": Conv(f"res{blk.name}_branch2a", blk.branch2a, blk.bn2a, stride=blk.stride),
                     "b": Conv(f"res{blk.name}_branch2b", blk.branch2b, blk.bn2b),
                     "c": Conv(f"res{blk.name}_branch2c", blk.branch2c, blk.bn2c),
                     "s": Conv(f"res{blk.name}_branch1", blk.branch1, blk.bn1, stride=blk.stride) if blk.project else None,
                     "stride": blk.stride}
                st.append(d)
            self.blocks.append(st)
        order = []
        for st in self.blocks[::-1]:
            for d in st[::-1]:
                order += [d["c"], d["b"], d["a"]] + ([d["s"]] if d["s"] is not None else [])
        order.append(self.stem)
        return order

    def _build_top(self):
        """FPN + heads in backward-completion order."""
        m = self.model
        f = m.fpn
        self.fpn = {n: Conv(n, getattr(f, n), stride=(2 if n in ("P6", "P7") else 1), same=n in ("P6", "P7"))
                    for n in f.LAYER_ORDER}
        self.reg_tower = [Conv(f"pyramid_regression_{i}", c) for i, c in enumerate(m.regression_submodel.tower)]
        self.reg_out = Conv("pyramid_regression", m.regression_submodel.out)
        self.cls_tower = [Conv(f"pyramid_classification_{i}", c) for i, c in enumerate(m.classification_submodel.tower)]
        self.cls_out = Conv("pyramid_classification", m.classification_submodel.out)
        order = [self.cls_out] + self.cls_tower[::-1] + [self.reg_out] + self.reg_tower[::-1]
        order += [self.fpn[n] for n in ("P7", "P6", "P3", "C3_reduced", "P4", "C4_reduced", "P5", "C5_reduced")]
        return order

    # hooks for engines that keep part of the model outside the conv table (engine/hybrid.py)
    def _alloc_extra(self, off: int, trainable: bool) -> int:
        return off

    def _bind_extra(self):
        pass

    def _alloc_params(self):
        dev = self.device
        off = 0
        for c in self.convs:  # trainable first so that [0, n_train) is the allreduce range
            if not c.trainable:
                continue
            c.w_off = off
            off = _align4(off + c.taps * c.cout * c.cin)
            if c.has_bias:
                c.b_off = off
                off = _align4(off + c.cout)
        off = self._alloc_extra(off, True)
        self.n_train = off
        # [0, cut) of G is final once the layer that ends there has run its weight gradient (parameters are laid
        # out in backward-completion order): where the overlapped allreduce may start sending
        self._cut_after = {c.name: (c.b_off + c.cout if c.has_bias else c.w_off + c.taps * c.cout * c.cin)
                           for c in self.convs if c.trainable}
        for c in self.convs:
            if c.trainable:
                continue
            c.w_off = off
            off = _align4(off + c.taps * c.cout * c.cin)
            if c.has_bias:
                c.b_off = off
                off = _align4(off + c.cout)
        off = self._alloc_extra(off, False)
        self.n_params = off
        pad = 64
        self.P = torch.zeros(off + pad, device=dev)
        self.M = torch.zeros(self.n_train + pad, device=dev)
        self.V = torch.zeros(self.n_train + pad, device=dev)
        self.G = torch.zeros(self.n_train + pad, device=dev)
        self.Gavg = self.G
        # frozen BN affine per conv
        nbn = sum(round8(c.cout) for c in self.convs if c.bn is not None)
        self.bn_scale = torch.ones(nbn, device=dev)
        self.bn_shift = torch.zeros(nbn, device=dev)
        o = 0
        for c in self.convs:
            if c.bn is not None:
                c.scale = self.bn_scale[o:o + c.cout]
                c.shift = self.bn_shift[o:o + c.cout]
                o += round8(c.cout)
            else:
                c.scale = None
                c.shift = None
        # bf16 working weights
        entries = []
        for c in self.convs:
            if c is self.stem:
                continue
            c.wf = torch.zeros(c.taps, c.cout, round8(c.cin), device=dev, dtype=BF16)
            need_d = not (self.freeze_backbone and not c.trainable)
            c.wd = torch.zeros(c.taps, c.cin, round8(c.cout), device=dev, dtype=BF16) if need_d else None
            entries.append(dict(src=c.wview(self.P), scale=c.scale, dst_f=c.wf, dst_d=c.wd, taps=c.taps, Cout=c.cout,
                                Cin=c.cin))
        self.wtable = native.WeightTable(entries)
        # small device scalars
        self.scal = torch.zeros(8, device=dev)           # [0] sumsq  [1] clip scale  [2] reg loss  [3] cls loss
        self.npos = torch.zeros(1, dtype=torch.int32, device=dev)
        self.ones64 = torch.ones(64, device=dev)
        self.zeros64 = torch.zeros(64, device=dev)
        self.lr_dev = torch.zeros(1, device=dev)          # bias-corrected Adam step size of the current step
        self.loss_out = torch.zeros(3, device=dev)        # [loss, regression_loss, classification_loss]
        self.loss_host = torch.zeros(3).pin_memory()      # the same triple, copied out right after the loss kernels
        self._comm_stream = torch.cuda.Stream(device=dev)  # gradient scatter, beside the backward pass
        self._pushed = 0                                  # elements of G already scattered in this step
        self._ev_front = torch.cuda.Event()               # forward + losses of the latest step are done
        self._ev_h2d = torch.cuda.Event()
        self._copy_stream = torch.cuda.Stream(device=dev)
        self._steps_issued = 0
        self.stem_wp = torch.zeros(64, 256, device=dev, dtype=BF16)
        self._bind_extra()

    # ------------------------------------------------------------------ parameter import / export
    @torch.no_grad()
    def load_from_model(self, model: RetinaNet | None = None):
        for c in self.convs:
            w = c.module.weight.detach().to(self.device, torch.float32)
            c.wview(self.P).copy_(w.permute(2, 3, 0, 1).reshape(c.taps, c.cout, c.cin))
            if c.has_bias:
                c.bview(self.P).copy_(c.module.bias.detach().to(self.device, torch.float32))
            if c.bn is not None:
                s, t = c.bn.scale_shift()
                c.scale.copy_(s.to(self.device))
                c.shift.copy_(t.to(self.device))
        self.__dict__.pop("_stem_idx", None)      # the packed-stem helper caches the BN scale
        self._refresh_weights()

    @torch.no_grad()
    def store_to_model(self, model: RetinaNet | None = None):
        for c in self.convs:
            w = c.wview(self.P).view(c.kh, c.kw, c.cout, c.cin).permute(2, 3, 0, 1)
            c.module.weight.data.copy_(w.to(c.module.weight.device))
            if c.has_bias:
                c.module.bias.data.copy_(c.bview(self.P).to(c.module.bias.device))
        return self.model

    def _refresh_weights(self):
        """fp32 masters -> bf16 working weights (BN scale folded): one table-driven kernel for every
        TMA-fed conv plus a gather for the stem's packed [64][256] filter."""
        self.wtable.run()
        s = self.stem
        if s is None:
            return
        if not hasattr(self, "_stem_idx"):
            # packed k = r*28 + s*4 + c  <-  master (tap = r*7 + s, co, ci); padding entries read a zero slot
            k = torch.arange(256, device=self.device)
            r, rem = k // 28, k % 28
            sx, c = rem // 4, rem % 4
            ok = (k < 196) & (c < 3)
            co = torch.arange(64, device=self.device).view(64, 1)
            idx = ((r * 7 + sx).view(1, 256) * 64 + co) * 3 + c.view(1, 256)
            self._stem_idx = torch.where(ok.view(1, 256), idx, torch.zeros_like(idx))
            self._stem_ok = ok.view(1, 256).float() * s.scale.view(64, 1)
        self.stem_wp.copy_(s.wview(self.P).reshape(-1)[self._stem_idx] * self._stem_ok)   # in place: address is stable

    def optimizer_state(self):
        """Adam slots as {keras layer name: {'kernel': (m, v), 'bias': (m, v)}} views + iterations."""
        return {"iterations": self.iterations, "lr": self.lr, "M": self.M, "V": self.V}

    # ------------------------------------------------------------------ per-shape plan
    def _plan_backbone(self, p, B, H, W, buf):
        """Backbone activations and gradient buffers; returns the (h, w) of C3, C4, C5."""
        dev = self.device
        p["img"] = torch.empty(B, H, W, 4, device=dev, dtype=BF16)   # 4th channel is zero padding
        h1, w1 = -(-H // 2), -(-W // 2)
        p["stem_raw_hw"] = (h1, w1)
        ph, pw = tf_same_pad(h1, 3, 2), tf_same_pad(w1, 3, 2)
        p["pool_pad"] = (ph[0], pw[0])
        h2, w2 = -(-h1 // 2), -(-w1 // 2)
        p["pool"] = buf(h2, w2, 64)
        p["stem_out"] = buf(h1, w1, 64)
        p["pool_idx"] = torch.empty(B, h2, w2, 64, device=dev, dtype=torch.uint8)
        acts = []
        hh, ww = h2, w2
        for st in self.blocks:
            sacts = []
            for d in st:
                s = d["stride"]
                ho, wo = -(-hh // s), -(-ww // s)
                f = d["a"].cout
                sacts.append({"a": buf(ho, wo, f), "b": buf(ho, wo, f), "o": buf(ho, wo, f * 4),
                              "s": buf(ho, wo, f * 4) if d["s"] is not None else None, "in_hw": (hh, ww),
                              "hw": (ho, wo)})
                hh, ww = ho, wo
            acts.append(sacts)
        p["acts"] = acts
        # backbone gradient buffers, sized per stage
        gb = []
        for si, st in enumerate(self.blocks):
            a0 = acts[si][0]
            f = st[0]["a"].cout
            ho, wo = a0["hw"]
            # gx = gradient w.r.t. the stage's input = the previous stage's output gradient: the same buffer
            gx = gb[si - 1]["g"] if si > 0 else buf(*a0["in_hw"], st[0]["a"].cin)
            assert tuple(gx.shape[1:]) == (*a0["in_hw"], st[0]["a"].cin)
            gb.append({"g": buf(ho, wo, f * 4), "ga": buf(ho, wo, f), "gb": buf(ho, wo, f), "gx": gx,
                       "tmp": buf(*a0["in_hw"], st[0]["a"].cin) if si == 0 else None})
        p["gb"] = gb
        p["dstem"] = torch.empty(B, h1, w1, 64, device=dev, dtype=BF16)
        return acts[1][-1]["hw"], acts[2][-1]["hw"], acts[3][-1]["hw"]

    def _plan(self, B, H, W):
        key = (B, H, W)
        if key in self._plans:
            self._plans[key] = self._plans.pop(key)   # most recently used last
            return self._plans[key]
        while len(self._plans) >= self.max_plans:
            self._plans.pop(next(iter(self._plans)))  # evict the least recently used plan
        dev = self.device
        p = {}

        def buf(h, w, c, zero=False):
            return (torch.zeros if zero else torch.empty)(B, h, w, c, device=dev, dtype=BF16)

        c3hw, c4hw, c5hw = self._plan_backbone(p, B, H, W, buf)
        p6hw = (-(-c5hw[0] // 2), -(-c5hw[1] // 2))
        p7hw = (-(-p6hw[0] // 2), -(-p6hw[1] // 2))
        lv = [c3hw, c4hw, c5hw, p6hw, p7hw]
        p["levels"] = lv
        fp = {"c5r": buf(*c5hw, 256), "m4": buf(*c4hw, 256), "m3": buf(*c3hw, 256), "r6": buf(*p6hw, 256)}
        fp["P"] = [buf(h, w, 256) for (h, w) in lv]
        p["fpn"] = fp
        K = self.num_classes
        p["cls_t"] = [[buf(h, w, 256) for (h, w) in lv] for _ in range(4)]
        p["reg_t"] = [[buf(h, w, 256) for (h, w) in lv] for _ in range(4)]
        p["cls_o"] = [buf(h, w, round8(9 * K), zero=True) for (h, w) in lv]
        p["reg_o"] = [buf(h, w, 40, zero=True) for (h, w) in lv]
        # gradient scratch
        p["dcls_o"] = [buf(h, w, round8(9 * K), zero=True) for (h, w) in lv]
        p["dreg_o"] = [buf(h, w, 40, zero=True) for (h, w) in lv]
        p["dt"] = [[buf(h, w, 256) for (h, w) in lv] for _ in range(2)]
        p["dP"] = [buf(h, w, 256) for (h, w) in lv]
        p["dfpn"] = {"m3": buf(*c3hw, 256), "m4": buf(*c4hw, 256), "c5r": buf(*c5hw, 256), "p6": buf(*p6hw, 256)}
        # anchors + targets
        p["anchors"] = native.anchors_for_shape_gpu(lv, dev)
        A = p["anchors"].shape[0]
        p["A"] = A
        p["state"] = torch.empty(B, A, dtype=torch.int8, device=dev)
        p["cls_id"] = torch.empty(B, A, dtype=torch.int16, device=dev)
        p["reg_t_tgt"] = torch.empty(B, A, 4, device=dev)
        p["a_off"] = [0]
        for (h, w) in lv:
            p["a_off"].append(p["a_off"][-1] + h * w * 9)
        self._plans[key] = p
        return p

    # ------------------------------------------------------------------ kernels wrappers
    def _fprop(self, c: Conv, x, out, relu=False, add=None):
        h, w = x.shape[1:3]
        ph, pw, _, _ = c.pads(h, w)
        shift = c.shift if c.bn is not None else (c.bview(self.P) if c.has_bias else None)
        native.conv_raw(0, [(x, out, add, None)], c.wf, c.cin, x.shape[-1], c.cout, out.shape[-1], c.kh, c.kw,
                        c.stride, ph, pw, shift=shift, relu=relu)

    def _fprop_group(self, c: Conv, xs, outs, relu=False):
        native.conv_raw(0, [(x, o, None, None) for x, o in zip(xs, outs)], c.wf, c.cin, xs[0].shape[-1], c.cout,
                        outs[0].shape[-1], c.kh, c.kw, 1, c.kh // 2, c.kw // 2, shift=c.bview(self.P), relu=relu)

    def _dgrad(self, c: Conv, dy, out, mask=None, add=None, add_after_mask=False, colsum=None):
        """Gradient w.r.t. the conv input.  out [N,H,W,cin].  ``colsum``: fp32 vector that receives the sum of ``out``
        over pixels -- the bias gradient of the layer whose dY this is -- computed by the kernel's store warp
        (with stride 2 every parity class must then have taps, which holds for 3x3)."""
        h, w = out.shape[1:3]
        ph, pw, _, _ = c.pads(h, w)
        if c.stride == 1:
            native.conv_raw(1, [(dy, out, add, mask, colsum)], c.wd, c.cout, dy.shape[-1], c.cin, out.shape[-1], c.kh,
                            c.kw, 1, ph, pw, add_after_mask=add_after_mask)
            return
        for py in range(2):
            for px in range(2):
                ok = any((py + ph - r) % 2 == 0 for r in range(c.kh)) and any((px + pw - s) % 2 == 0 for s in range(c.kw))
                assert ok or colsum is None
                if ok:
                    native.conv_raw(1, [(dy, out, add, mask, colsum)], c.wd, c.cout, dy.shape[-1], c.cin, out.shape[-1],
                                    c.kh, c.kw, 2, ph, pw, par=(py, px), add_after_mask=add_after_mask)

    def _dgrad_group(self, c: Conv, dys, outs, masks=None, adds=None, colsums=None):
        probs = [(dy, o, None if adds is None else adds[i], None if masks is None else masks[i],
                  None if colsums is None else colsums[i]) for i, (dy, o) in enumerate(zip(dys, outs))]
        native.conv_raw(1, probs, c.wd, c.cout, dys[0].shape[-1], c.cin, outs[0].shape[-1], c.kh, c.kw, 1, c.kh // 2,
                        c.kw // 2)

    def _wgrad(self, c: Conv, xs, dys, dbias=True):
        """``dbias=False``: the bias gradient was already produced by the kernel that wrote ``dys`` (colsum)."""
        if not c.trainable:
            return
        h, w = xs[0].shape[1:3]
        ph, pw, _, _ = c.pads(h, w)
        native.wgrad_raw(list(zip(xs, dys)), c.wview(self.G), c.cin, xs[0].shape[-1], c.cout, dys[0].shape[-1], c.kh,
                         c.kw, c.stride, ph, pw, rowscale=c.scale,
                         dbias=c.bview(self.G) if (c.has_bias and (dbias or not self._fused_bias)) else None)

    def _bias_grad(self, c: Conv):
        """Target of the fused column sums, or None (then ``_wgrad(..., dbias=False)`` must not be used: see
        ``_fused_bias``)."""
        return c.bview(self.G) if (c.has_bias and c.trainable and self._fused_bias) else None

    # ------------------------------------------------------------------ forward
    def _forward(self, p, images_u8, img_hw=None):
        p["C"] = self._forward_backbone(p, images_u8, img_hw)
        self._forward_top(p, *p["C"])

    def _forward_backbone(self, p, images_u8, img_hw=None):
        """Pixels -> (C3, C4, C5) NHWC bf16."""
        native.preprocess(images_u8, p["img"], CAFFE_MEAN_BGR, img_hw)
        native.stem_conv(p["img"], self.stem_wp, self.stem.shift, p["stem_out"])     # conv + BN + ReLU
        native.bn_relu_maxpool_fwd(p["stem_out"], self.ones64, self.zeros64, p["pool"], p["pool_idx"], *p["pool_pad"])
        self._mark("fwd_stem")
        x = p["pool"]
        for si, st in enumerate(self.blocks):
            for bi, d in enumerate(st):
                a = p["acts"][si][bi]
                self._fprop(d["a"], x, a["a"], relu=True)
                self._fprop(d["b"], a["a"], a["b"], relu=True)
                if d["s"] is not None:
                    self._fprop(d["s"], x, a["s"])
                    self._fprop(d["c"], a["b"], a["o"], relu=True, add=a["s"])
                else:
                    self._fprop(d["c"], a["b"], a["o"], relu=True, add=x)
                a["x"] = x
                x = a["o"]
        self._mark("fwd_backbone")
        return p["acts"][1][-1]["o"], p["acts"][2][-1]["o"], p["acts"][3][-1]["o"]

    def _forward_top(self, p, c3, c4, c5):
        """FPN + the two shared heads over five levels."""
        fp, f = p["fpn"], self.fpn
        P3, P4, P5, P6, P7 = fp["P"]
        self._fprop(f["C5_reduced"], c5, fp["c5r"])
        self._fprop(f["P5"], fp["c5r"], P5)
        self._fprop(f["C4_reduced"], c4, fp["m4"])
        native.upsample_add(fp["m4"], fp["c5r"], fp["m4"])
        self._fprop(f["P4"], fp["m4"], P4)
        self._fprop(f["C3_reduced"], c3, fp["m3"])
        native.upsample_add(fp["m3"], fp["m4"], fp["m3"])
        self._fprop(f["P3"], fp["m3"], P3)
        self._fprop(f["P6"], c5, P6)
        native.relu(P6, fp["r6"])
        self._fprop(f["P7"], fp["r6"], P7)
        self._mark("fwd_fpn")
        for tower, outc, ts, outs in ((self.cls_tower, self.cls_out, p["cls_t"], p["cls_o"]),
                                      (self.reg_tower, self.reg_out, p["reg_t"], p["reg_o"])):
            xs = fp["P"]
            for i, c in enumerate(tower):
                self._fprop_group(c, xs, ts[i], relu=True)
                xs = ts[i]
            self._fprop_group(outc, xs, outs)

    # ------------------------------------------------------------------ losses (fused fwd+bwd)
    def _losses(self, p, gt, gt_count, img_hw):
        self.npos.zero_()
        self.scal.zero_()
        native.assign_targets(p["anchors"], gt, gt_count, img_hw, p["state"], p["cls_id"], p["reg_t_tgt"], self.npos)
        A = p["A"]
        for l in range(5):
            native.focal_loss(p["cls_o"][l], p["dcls_o"][l], self.num_classes, p["state"], p["cls_id"], A,
                              p["a_off"][l], self.npos, self.scal[3:4])
            native.smooth_l1(p["reg_o"][l], p["dreg_o"][l], p["state"], p["reg_t_tgt"], A, p["a_off"][l], self.npos,
                             self.scal[2:3])

    # ------------------------------------------------------------------ overlapped gradient scatter
    def _overlap(self) -> bool:
        return self.comm is not None and self.comm.world > 1 and self.comm.can_stage()

    def _push_ready(self, upto: int):
        """G[:upto] is final: scatter the new part to its owner ranks on the side stream (raw, un-clipped: the
        clip factor of reference train.py:103-104 is applied by the owners once every rank knows its norm)."""
        if not self._overlap():
            return
        upto = min(upto, self.n_train) // 4 * 4
        if upto <= self._pushed:
            return
        main = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(main)
        self._comm_stream.wait_event(ev)
        if self._comm_prof is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(self._comm_stream)
        self.comm.scatter(self.n_train, self._pushed, upto, stream=self._comm_stream)
        if self._comm_prof is not None:
            e1.record(self._comm_stream)
            self._comm_prof.append((self._pushed, upto, e0, e1))
        self._pushed = upto

    def _push_after(self, conv):
        if conv.trainable:
            self._push_ready(self._cut_after[conv.name])

    # ------------------------------------------------------------------ backward
    def _backward(self, p):
        self._pushed = 0
        fp, f = p["fpn"], self.fpn
        feats = fp["P"]
        dP = p["dP"]
        first = True
        for tower, outc, ts, douts in ((self.cls_tower, self.cls_out, p["cls_t"], p["dcls_o"]),
                                       (self.reg_tower, self.reg_out, p["reg_t"], p["dreg_o"])):
            # Bias gradients: the dgrad that PRODUCES a layer's dY also sums it over pixels (store-warp column sums),
            # so only the two output convs, whose dY comes from the loss kernels, and C4_reduced, whose dY is
            # touched up by the upsample backward, still run the column-sum kernel.
            self._wgrad(outc, ts[3], douts)
            cur = p["dt"][0]
            self._dgrad_group(outc, douts, cur, masks=ts[3], colsums=[self._bias_grad(tower[3])] * 5)
            for i in (3, 2, 1):
                self._wgrad(tower[i], ts[i - 1], cur, dbias=False)
                nxt = p["dt"][1] if cur is p["dt"][0] else p["dt"][0]
                self._dgrad_group(tower[i], cur, nxt, masks=ts[i - 1], colsums=[self._bias_grad(tower[i - 1])] * 5)
                cur = nxt
            self._wgrad(tower[0], feats, cur, dbias=False)
            pbias = None if first else [self._bias_grad(f[n]) for n in ("P3", "P4", "P5", "P6", "P7")]
            if pbias is not None:
                pbias[3] = None        # P6's dY also gets P7's gradient below
            self._dgrad_group(tower[0], cur, dP, adds=None if first else dP, colsums=pbias)
            first = False
        self._push_after(self.reg_tower[0])
        self._mark("bwd_heads")
        dP3, dP4, dP5, dP6, dP7 = dP
        df = p["dfpn"]
        c3, c4, c5 = p["C"]
        P6 = feats[3]
        # P7 = conv(relu(P6)); P6 also feeds the heads directly
        self._wgrad(f["P7"], [fp["r6"]], [dP7], dbias=False)
        self._dgrad(f["P7"], dP7, df["p6"], mask=P6, add=dP6, add_after_mask=True, colsum=self._bias_grad(f["P6"]))
        self._wgrad(f["P6"], [c5], [df["p6"]], dbias=False)
        # P3 branch
        self._wgrad(f["P3"], [fp["m3"]], [dP3], dbias=False)
        self._dgrad(f["P3"], dP3, df["m3"], colsum=self._bias_grad(f["C3_reduced"]))
        self._wgrad(f["C3_reduced"], [c3], [df["m3"]], dbias=False)
        # P4 branch
        self._wgrad(f["P4"], [fp["m4"]], [dP4], dbias=False)
        self._dgrad(f["P4"], dP4, df["m4"])
        native.upsample_bwd_add(df["m3"], df["m4"])
        self._wgrad(f["C4_reduced"], [c4], [df["m4"]])
        # P5 branch
        self._wgrad(f["P5"], [fp["c5r"]], [dP5], dbias=False)
        self._dgrad(f["P5"], dP5, df["c5r"])
        native.upsample_bwd_add(df["m4"], df["c5r"])
        self._wgrad(f["C5_reduced"], [c5], [df["c5r"]])
        self._push_after(f["C5_reduced"])
        self._mark("bwd_fpn")
        if not self.freeze_backbone:
            self._backward_backbone(p)

    def _backward_backbone(self, p):
        fp, f, df = p["fpn"], self.fpn, p["dfpn"]
        c3, c4, c5 = p["C"]
        gb = p["gb"]
        # gradient w.r.t. C5 / C4 / C3 from the FPN (masked by the ReLU of the block output)
        g5 = gb[3]["g"]
        self._dgrad(f["C5_reduced"], df["c5r"], g5)
        self._dgrad(f["P6"], df["p6"], g5, mask=c5, add=g5)
        extra = {2: (f["C4_reduced"], df["m4"]), 1: (f["C3_reduced"], df["m3"])}
        g = g5
        for si in (3, 2, 1, 0):
            st = self.blocks[si]
            acts = p["acts"][si]
            b = gb[si]
            for bi in range(len(st) - 1, -1, -1):
                d, a = st[bi], acts[bi]
                x = a["x"]
                self._wgrad(d["c"], [a["b"]], [g])
                self._dgrad(d["c"], g, b["gb"], mask=a["b"])
                self._wgrad(d["b"], [a["a"]], [b["gb"]])
                self._dgrad(d["b"], b["gb"], b["ga"], mask=a["a"])
                self._wgrad(d["a"], [x], [b["ga"]])
                if d["s"] is None:
                    self._dgrad(d["a"], b["ga"], g, mask=x, add=g)  # in place: g <- mask(dgrad + g)
                else:
                    self._wgrad(d["s"], [x], [g])
                    gx = b["gx"]
                    if d["stride"] == 1:
                        self._dgrad(d["s"], g, b["tmp"])
                        self._dgrad(d["a"], b["ga"], gx, mask=x, add=b["tmp"])
                    else:
                        # Stride-2 block: the 1x1/2 convs reach only the even-even pixels of x.  Every contribution is
                        # masked by x's ReLU inside its own kernel (the mask distributes over the sum), so no
                        # separate add / mask / copy passes: the FPN lateral's dgrad (all pixels) or a zero fill
                        # comes first, then the two strided dgrads accumulate in place.
                        if (si - 1) in extra:
                            lat, dlat = extra[si - 1]
                            self._dgrad(lat, dlat, gx, mask=x)
                        else:
                            gx.zero_()
                        self._dgrad(d["s"], g, gx, mask=x, add=gx, add_after_mask=True)
                        self._dgrad(d["a"], b["ga"], gx, mask=x, add=gx, add_after_mask=True)
                    g_next = gx
            g = g_next      # stage si-1's output gradient (gb[si-1]["g"] is the same buffer as this stage's gx)
            if si >= 2:     # stages 5, 4 (98 % of the backbone's parameters); the small rest goes with the stem
                self._push_after(st[0]["s"])
        self._mark("bwd_backbone")
        # stem: max-pool + ReLU + BN backward, then the native weight-gradient kernel for the 7x7x3 filter
        native.bn_relu_maxpool_bwd(p["pool_idx"], self.ones64, g, p["dstem"], *p["pool_pad"])
        if self.stem.trainable:
            native.stem_wgrad(p["img"], p["dstem"], self.stem.scale, self.stem.wview(self.G))

    # ------------------------------------------------------------------ optimizer
    def _attach_comm(self):
        """Move the gradient bucket into peer-mapped memory, allocate the reduced copy and make every
        rank start from rank 0's state (Horovod BroadcastGlobalVariablesCallback(0), train.py:111)."""
        self.G = self.comm.symmetric_bucket(self.n_train + 64)
        self.Gavg = self.comm.result[:self.G.numel()]   # peer-writable: the allreduce pushes results into it
        self._bind_extra()
        self.broadcast_state(0)

    def broadcast_state(self, root: int = 0):
        if self.comm is None or self.comm.world == 1:
            return
        for t in (self.P, self.M, self.V, self.bn_scale, self.bn_shift):
            self.comm.broadcast_(t, root)
        it = torch.tensor([float(self.iterations), float(self.lr), 0.0, 0.0], device=self.device)
        self.comm.broadcast_(it, root)
        self.iterations, self.lr = int(it[0].item()), float(it[1].item())
        self.__dict__.pop("_stem_idx", None)      # the packed-stem helper caches the (just replaced) BN scale
        self._refresh_weights()

    def check_comm(self):
        """Raise if a gradient collective timed out since the last check (a peer rank is gone).  Synchronises."""
        if self.comm is not None and hasattr(self.comm, "check"):
            self.comm.check()

    def _begin_step(self):
        """Host-side per-step scalars, written to device memory OUTSIDE any captured graph."""
        self.iterations += 1
        self._steps_issued += 1
        t = self.iterations
        self.lr_dev.fill_(self.lr * math.sqrt(1.0 - self.beta2 ** t) / (1.0 - self.beta1 ** t))

    def _optimizer_step(self):
        n = self.n_train
        if self.comm is not None and self.comm.world > 1:
            with self._phase("allreduce"):
                native.sumsq(self.G[:n], self.scal[0:1])
                # the clip factor lives in peer-visible memory: peers apply it while reading my bucket
                native.clip_scale(self.scal[0:1], self.clipnorm, self.comm.prescale_local)
                if self._overlap():
                    # everything up to the last cut is already at its owner; send the tail, then reduce + gather
                    self._push_ready(n)
                    ev = torch.cuda.Event()
                    ev.record(self._comm_stream)
                    torch.cuda.current_stream(self.device).wait_event(ev)
                    self.comm.allreduce_staged_avg(n, prescale=True)
                else:
                    self.comm.allreduce_avg(self.G, self.Gavg, n, prescale=True)
            with self._phase("optimizer"):
                self._adam_and_refresh(distributed=True)
            return
        with self._phase("optimizer"):
            native.sumsq(self.G[:n], self.scal[0:1])
            self._adam_and_refresh(distributed=False)

    def _adam_and_refresh(self, distributed: bool):
        n = self.n_train
        if distributed:
            native.adam(self.P[:n], self.Gavg[:n], self.M[:n], self.V[:n], None, self.lr_dev, self.beta1, self.beta2,
                        self.eps)
        else:
            native.clip_scale(self.scal[0:1], self.clipnorm, self.scal[1:2])
            native.adam(self.P[:n], self.G[:n], self.M[:n], self.V[:n], self.scal[1:2], self.lr_dev, self.beta1,
                        self.beta2, self.eps)
        self._refresh_weights()

    def _mark(self, name):
        if self._prof is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._prof.append((name, ev))

    class _Phase:
        """NVTX range (B200_NVTX=1; shows in nsys / ncu timelines) + profile mark around one phase of the step."""

        def __init__(self, trainer, name):
            self.t, self.name = trainer, name

        def __enter__(self):
            if self.t._nvtx:
                torch.cuda.nvtx.range_push(self.name)

        def __exit__(self, *exc):
            if self.t._nvtx:
                torch.cuda.nvtx.range_pop()
            self.t._mark(self.name)
            return False

    def _phase(self, name):
        return NativeTrainer._Phase(self, name)

    def trace_step(self, *batch, tag: str = ""):
        """One eager step timed with CUDA events per phase -- on the main stream AND on the gradient-scatter stream --
        written into the runtime's Chrome-trace timeline (``--timeline`` / HOROVOD_TIMELINE) as GPU_<PHASE> events:
        the overlapped allreduce (GRAD_SCATTER pieces beside BACKWARD, ALLREDUCE after it) becomes visible next to
        the host runtime's phases.  Returns {phase: ms}."""
        from .. import parallel as hvd
        torch.cuda.synchronize(self.device)
        t_anchor = hvd.timeline_now_ns() if hvd.timeline_active() else 0
        self._comm_prof = []
        out = self.profile_step(*batch, keep=True)
        base = self._prof_events[0][1]
        if hvd.timeline_active():
            for (n0, e0), (n1, e1) in zip(self._prof_events[:-1], self._prof_events[1:]):
                hvd.timeline_event("GPU_" + n1.upper(), tag or "step", t_anchor + int(base.elapsed_time(e0) * 1e6),
                                   t_anchor + int(base.elapsed_time(e1) * 1e6))
            for (lo, hi, e0, e1) in self._comm_prof:
                hvd.timeline_event("GPU_GRAD_SCATTER", f"G[{lo}:{hi}]", t_anchor + int(base.elapsed_time(e0) * 1e6),
                                   t_anchor + int(base.elapsed_time(e1) * 1e6))
        out["grad_scatter_pieces"] = [(lo, hi, round(e0.elapsed_time(e1), 4)) for lo, hi, e0, e1 in self._comm_prof]
        self._comm_prof = None
        self._prof_events = None
        return out

    def profile_step(self, *batch, keep=False):
        """Run one step with CUDA events between phases; returns {phase: ms} (device time)."""
        self._prof = []
        self._mark("start")
        self.train_step(*batch)
        torch.cuda.synchronize()
        out = {}
        for (n0, e0), (n1, e1) in zip(self._prof[:-1], self._prof[1:]):
            out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1)
        if keep:
            self._prof_events = self._prof
        self._prof = None
        return out

    def _pad_bgr(self):
        """uint8 colour that preprocesses to (about) zero, the value the reference pads batches with."""
        return (104, 117, 124) if getattr(self.model, "preprocess_mode", "caffe") == "caffe" else (127, 127, 127)

    def _bucket(self, images_u8):
        """Pad H, W up to a multiple of ``shape_bucket`` with the mean colour (zero after preprocessing).
        Anchors whose centre falls in the padding are ignored through ``img_hw`` exactly as for the
        reference's batch padding."""
        b = self.shape_bucket
        H, W = images_u8.shape[1:3]
        Hp, Wp = -(-H // b) * b, -(-W // b) * b
        if (Hp, Wp) == (H, W):
            return images_u8
        out = torch.empty(images_u8.shape[0], Hp, Wp, 3, dtype=torch.uint8, device=images_u8.device)
        out[..., 0], out[..., 1], out[..., 2] = self._pad_bgr()
        out[:, :H, :W] = images_u8
        return out

    # ------------------------------------------------------------------ public API
    def train_step(self, images_u8, gt, gt_count, img_hw):
        """One optimisation step.  Inputs may live in (pinned) host memory or on the device:
        images uint8 [B,H,W,3] BGR, gt fp32 [B,G,5] (x1,y1,x2,y2,label), gt_count int32 [B],
        img_hw int32 [B,2] un-padded sizes.  Returns a device tensor [loss, regression_loss,
        classification_loss] (the Keras metric names of reference train.py:99-102)."""
        return self.train_step_async(images_u8, gt, gt_count, img_hw).loss

    def train_step_async(self, images_u8, gt, gt_count, img_hw) -> "StepHandle":
        """Same step, returning a handle: ``handle.loss`` is the device tensor ``train_step`` returns,
        ``handle.loss_cpu()`` the same three numbers on the host, available as soon as the forward pass and
        the losses are done -- backward, allreduce and Adam keep running, and the next call's host-to-device
        copy of its inputs overlaps them (it only waits for this step to have consumed its inputs)."""
        dev = self.device
        B, H, W = images_u8.shape[:3]
        b = self.shape_bucket
        p = self._plan(B, -(-H // b) * b, -(-W // b) * b)
        self._begin_step()
        main = torch.cuda.current_stream(dev)
        if "graph" in p and self._graph_usable(p, gt):
            # steady state: inputs go straight (H2D when they are on the host) into the graph's static buffers
            self._load_static(p, images_u8, gt, gt_count, img_hw)
            ga, gb = p["graph"]
            ga.replay()
            self._ev_front.record(main)
            gb.replay()
            native.add_launch_count(p["graph_kernels"])
            return StepHandle(self)
        images_u8 = images_u8.to(dev, non_blocking=True)
        gt = gt.to(dev, torch.float32, non_blocking=True)
        gt_count = gt_count.to(dev, torch.int32, non_blocking=True)
        img_hw = img_hw.to(dev, torch.int32, non_blocking=True)
        images_u8 = self._bucket(images_u8)
        if self._graph_usable(p, gt):
            self._graph_step(p, images_u8, gt, gt_count, img_hw)
        else:
            self._step_body(p, images_u8, gt, gt_count, img_hw)
        return StepHandle(self)

    def _step_front(self, p, images_u8, gt, gt_count, img_hw):
        """Forward + losses; the loss triple is final here and goes to pinned host memory."""
        with self._phase("h2d+zero"):
            self.G.zero_()
        with self._phase("forward"):
            self._forward(p, images_u8, img_hw)
        with self._phase("losses"):
            self._losses(p, gt, gt_count, img_hw)
            torch.stack([self.scal[2] + self.scal[3], self.scal[2], self.scal[3]], out=self.loss_out)
            self.loss_host.copy_(self.loss_out, non_blocking=True)

    def _step_back(self, p):
        with self._phase("backward"):
            self._backward(p)
        self._optimizer_step()

    def _step_body(self, p, images_u8, gt, gt_count, img_hw):
        """A whole step with plain launches, in stream order."""
        self._step_front(p, images_u8, gt, gt_count, img_hw)
        self._ev_front.record(torch.cuda.current_stream(self.device))
        self._step_back(p)

    # ------------------------------------------------------------------ whole-step CUDA graphs
    def _graph_usable(self, p, gt) -> bool:
        if not self.use_graph or self._prof is not None or p.get("graph_failed"):
            return False
        if self.comm is not None and getattr(self.comm, "external_sync", False):
            return False                     # in-process towers order their streams with host barriers
        return gt.shape[1] <= self.graph_gt_cap

    def _load_static(self, p, images_u8, gt, gt_count, img_hw):
        """Inputs -> the graphs' static buffers.  Host tensors are copied on a side stream that only waits for the
        PREVIOUS step's front half (which consumed the buffers), so the transfer overlaps that step's backward."""
        main = torch.cuda.current_stream(self.device)
        side = self._copy_stream if not images_u8.is_cuda else None
        if side is not None:
            side.wait_event(self._ev_front)
        with torch.cuda.stream(side if side is not None else main):
            dst = p["in_img"]
            H, W = images_u8.shape[1:3]
            if (H, W) != tuple(dst.shape[1:3]):          # bucket padding: colour that preprocesses to zero
                dst[..., 0], dst[..., 1], dst[..., 2] = self._pad_bgr()
            dst[:, :H, :W].copy_(images_u8, non_blocking=True)
            p["in_gt"][:, :gt.shape[1]].copy_(gt, non_blocking=True)
            p["in_cnt"].copy_(gt_count, non_blocking=True)
            p["in_hw"].copy_(img_hw, non_blocking=True)
        if side is not None:
            self._ev_h2d.record(side)
            main.wait_event(self._ev_h2d)

    def _graph_step(self, p, images_u8, gt, gt_count, img_hw):
        """Warm-up (eager) steps of a plan, then capture + first replay.  Inputs are on the device, bucketed."""
        p["graph_warm"] = p.get("graph_warm", 0) + 1
        if p["graph_warm"] <= self.graph_warmup:
            self._step_body(p, images_u8, gt, gt_count, img_hw)
            return
        B = images_u8.shape[0]
        p["in_img"] = torch.empty_like(images_u8)
        p["in_gt"] = torch.zeros(B, self.graph_gt_cap, 5, device=self.device)
        p["in_cnt"] = torch.zeros(B, dtype=torch.int32, device=self.device)
        p["in_hw"] = torch.zeros(B, 2, dtype=torch.int32, device=self.device)
        self._load_static(p, images_u8, gt, gt_count, img_hw)
        try:
            torch.cuda.synchronize(self.device)
            k0 = native.launch_count()
            ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            # thread_local: prefetch workers keep allocating / launching on their own streams while this thread captures
            with torch.cuda.graph(ga, capture_error_mode="thread_local"):
                self._step_front(p, p["in_img"], p["in_gt"], p["in_cnt"], p["in_hw"])
            with torch.cuda.graph(gb, capture_error_mode="thread_local"):
                self._step_back(p)
            p["graph_kernels"] = native.launch_count() - k0     # capture counted them once: that is the first replay
        except Exception as e:  # noqa: BLE001 - capture is an optimisation: fall back to eager launches
            warnings.warn(f"CUDA-graph capture of the training step failed ({e}); continuing with eager launches")
            p["graph_failed"] = True
            self._step_body(p, images_u8, gt, gt_count, img_hw)
            return
        p["graph"] = (ga, gb)
        ga.replay()
        self._ev_front.record(torch.cuda.current_stream(self.device))
        gb.replay()

    @torch.no_grad()
    def forward_loss(self, images_u8, gt, gt_count, img_hw):
        """Forward + losses + backward WITHOUT the optimizer (testing / gradient checks)."""
        dev = self.device
        images_u8 = self._bucket(images_u8.to(dev))
        B, H, W, _ = images_u8.shape
        p = self._plan(B, H, W)
        self.G.zero_()
        img_hw = img_hw.to(dev, torch.int32)
        self._forward(p, images_u8, img_hw)
        self._losses(p, gt.to(dev, torch.float32), gt_count.to(dev, torch.int32), img_hw)
        self._backward(p)
        return torch.stack([self.scal[2] + self.scal[3], self.scal[2], self.scal[3]]), p

    @torch.no_grad()
    def predict(self, images_u8, score_threshold=0.05, max_detections=300, nms_threshold=0.5):
        """Prediction graph (reference train.py:95): boxes [B,300,4], scores [B,300], labels [B,300]."""
        images_u8 = images_u8.to(self.device)
        H0, W0 = images_u8.shape[1:3]
        images_u8 = self._bucket(images_u8)
        B, H, W, _ = images_u8.shape
        p = self._plan(B, H, W)
        self._forward(p, images_u8)
        return native.filter_detections_levels(p["cls_o"], p["reg_o"], self.num_classes, p["anchors"], (H0, W0),
                                               score_threshold, nms_threshold, max_detections)

    def gradients_by_layer(self):
        """{keras layer name: (dW OIHW fp32, db or None)} from the flat gradient buffer (tests)."""
        out = {}
        for c in self.convs:
            if not c.trainable:
                continue
            w = c.wview(self.G).view(c.kh, c.kw, c.cout, c.cin).permute(2, 3, 0, 1).contiguous()
            out[c.name] = (w, c.bview(self.G).clone() if c.has_bias else None)
        return out

    # ------------------------------------------------------------------ checkpoint surface
    def optimizer_config(self) -> dict:
        return {"lr": self.lr, "beta_1": self.beta1, "beta_2": self.beta2, "epsilon": self.eps, "decay": 0.0,
                "amsgrad": False, "clipnorm": self.clipnorm}

    def optimizer_slots(self) -> dict:
        """{keras weight name: (m, v)} in the torch layout (OIHW kernels), on the host."""
        out = {}
        for c in self.convs:
            if not c.trainable:
                continue
            m = c.wview(self.M).view(c.kh, c.kw, c.cout, c.cin).permute(2, 3, 0, 1).contiguous().cpu()
            v = c.wview(self.V).view(c.kh, c.kw, c.cout, c.cin).permute(2, 3, 0, 1).contiguous().cpu()
            out[f"{c.name}/kernel:0"] = (m, v)
            if c.has_bias:
                out[f"{c.name}/bias:0"] = (c.bview(self.M).cpu().clone(), c.bview(self.V).cpu().clone())
        return out

    @torch.no_grad()
    def load_optimizer_state(self, state: dict):
        cfg = state.get("optimizer_config") or {}
        self.lr = float(cfg.get("lr", self.lr))
        self.clipnorm = float(cfg.get("clipnorm", self.clipnorm) or 0.0)
        self.iterations = int(state.get("iterations", 0))
        slots = state.get("slots", {})
        for c in self.convs:
            if not c.trainable:
                continue
            k = slots.get(f"{c.name}/kernel:0")
            if k is not None:
                for flat, t in ((self.M, k[0]), (self.V, k[1])):
                    c.wview(flat).copy_(t.to(self.device, torch.float32).permute(2, 3, 0, 1).reshape(c.taps, c.cout, c.cin))
            b = slots.get(f"{c.name}/bias:0")
            if b is not None and c.has_bias:
                c.bview(self.M).copy_(b[0].to(self.device))
                c.bview(self.V).copy_(b[1].to(self.device))

    def save(self, path: str, epoch: int | None = None, weights_only: bool = False):
        """Keras-layout ``.h5`` snapshot (ModelCheckpoint, reference ``train.py:112-114``)."""
        from ..utils import keras_h5
        self.store_to_model()
        if weights_only:
            keras_h5.save_weights(self.model, path)
        else:
            keras_h5.save_model(self.model, path, optimizer_config=self.optimizer_config(),
                                optimizer_slots=self.optimizer_slots(), iterations=self.iterations, epoch=epoch)
