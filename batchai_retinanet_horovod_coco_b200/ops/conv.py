"""Convenience layer over the tcgen05 convolution kernels (``csrc/conv_sm100.cu``,
``csrc/wgrad_sm100.cu``): shape arithmetic, weight packing, output allocation.

Conventions
-----------
* activations: NHWC bf16 tensors ``[N, H, W, ld]`` (``ld`` >= channels, multiple of 8);
* master weights (engine layout): fp32 ``[KH*KW, Cout, Cin]``;
* working weights: bf16 ``wf [taps, Cout, ldf]`` for fprop and ``wd [taps, Cin, ldd]`` for dgrad
  (``ld*`` = channels rounded up to 8), with the frozen-BN scale folded in;
* padding is "padding before" per dim; TensorFlow ``same`` for stride 2 may be asymmetric
  (``same_pad_before``), which the kernels get for free from TMA out-of-bounds fill.

Replaces cuDNN conv fwd/dgrad/wgrad as used by the reference through TensorFlow
(``train.py:98-105,444``).
"""

from __future__ import annotations

import torch

from . import native


def round8(n: int) -> int:
    return (n + 7) // 8 * 8


def out_size(n: int, k: int, s: int, pad_total: int) -> int:
    return (n + pad_total - k) // s + 1


def pack_master(w_oihw: torch.Tensor) -> torch.Tensor:
    """torch OIHW fp32 -> engine master layout [KH*KW, Cout, Cin] fp32."""
    co, ci, kh, kw = w_oihw.shape
    return w_oihw.permute(2, 3, 0, 1).reshape(kh * kw, co, ci).contiguous().float()


def unpack_master(w_tci: torch.Tensor, kh: int, kw: int) -> torch.Tensor:
    t, co, ci = w_tci.shape
    return w_tci.reshape(kh, kw, co, ci).permute(2, 3, 0, 1).contiguous()


def pack_working(master: torch.Tensor, scale: torch.Tensor | None = None):
    """Plain-PyTorch version of ``b200_prepare_weights`` (oracle + small-model helper)."""
    t, co, ci = master.shape
    w = master if scale is None else master * scale.view(1, co, 1)
    wf = torch.zeros(t, co, round8(ci), device=master.device, dtype=torch.bfloat16)
    wf[:, :, :ci] = w.to(torch.bfloat16)
    wd = torch.zeros(t, ci, round8(co), device=master.device, dtype=torch.bfloat16)
    wd[:, :, :co] = w.transpose(1, 2).to(torch.bfloat16)
    return wf, wd


def fprop(x: torch.Tensor, wf: torch.Tensor, cin: int, cout: int, kh: int, kw: int, stride: int = 1,
          pad=(0, 0), pad_total=None, shift=None, relu=False, add=None, out=None, ldo=None, out_fp32=False):
    """x NHWC [N,H,W,ldi] -> out NHWC [N,OH,OW,ldo]."""
    n, h, w, ldi = x.shape
    if pad_total is None:
        pad_total = (2 * pad[0], 2 * pad[1])
    oh, ow = out_size(h, kh, stride, pad_total[0]), out_size(w, kw, stride, pad_total[1])
    ldo = ldo or round8(cout)
    if out is None:
        alloc = torch.zeros if ldo != cout else torch.empty
        out = alloc(n, oh, ow, ldo, device=x.device, dtype=torch.float32 if out_fp32 else torch.bfloat16)
    native.conv_raw(0, [(x, out, add, None)], wf, cin, ldi, cout, ldo, kh, kw, stride, pad[0], pad[1],
                    shift=shift, relu=relu, out_fp32=out_fp32)
    return out


def fprop_group(xs, wf, cin, cout, kh, kw, pad=(1, 1), shift=None, relu=False, outs=None, ldo=None):
    """Same-weight stride-1 conv applied to several tensors (pyramid levels) in ONE launch."""
    ldo = ldo or round8(cout)
    if outs is None:
        alloc = torch.zeros if ldo != cout else torch.empty
        outs = [alloc(x.shape[0], x.shape[1] + 2 * pad[0] - kh + 1, x.shape[2] + 2 * pad[1] - kw + 1, ldo,
                      device=x.device, dtype=torch.bfloat16) for x in xs]
    native.conv_raw(0, [(x, o, None, None) for x, o in zip(xs, outs)], wf, cin, xs[0].shape[-1], cout, ldo, kh, kw,
                    1, pad[0], pad[1], shift=shift, relu=relu)
    return outs


def dgrad(dy: torch.Tensor, wd: torch.Tensor, cin: int, cout: int, kh: int, kw: int, in_hw, stride: int = 1,
          pad=(0, 0), mask=None, add=None, out=None, ldo=None):
    """dy NHWC [N,OH,OW,ldy] -> dx NHWC [N,H,W,ldo].  ``mask``: dx = (mask > 0) ? dx : 0 (ReLU backward of
    the producer of x); ``add``: dx += add (gradient accumulation), applied before the mask."""
    n = dy.shape[0]
    h, w = in_hw
    ldo = ldo or round8(cin)
    if out is None:
        need_zero = stride == 2 or ldo != cin
        out = (torch.zeros if need_zero else torch.empty)(n, h, w, ldo, device=dy.device, dtype=torch.bfloat16)
    if stride == 1:
        native.conv_raw(1, [(dy, out, add, mask)], wd, cout, dy.shape[-1], cin, ldo, kh, kw, 1, pad[0], pad[1])
    else:
        for py in range(2):
            for px in range(2):
                has_taps = any((py + pad[0] - r) % 2 == 0 for r in range(kh)) and \
                    any((px + pad[1] - c) % 2 == 0 for c in range(kw))
                if not has_taps:
                    # no filter tap lands on this parity class: gradient is just (add) * mask
                    if add is not None:
                        sub = add[:, py::2, px::2, :]
                        if mask is not None:
                            sub = sub * (mask[:, py::2, px::2, :] > 0)
                        out[:, py::2, px::2, :] = sub
                    continue
                native.conv_raw(1, [(dy, out, add, mask)], wd, cout, dy.shape[-1], cin, ldo, kh, kw, 2, pad[0], pad[1],
                                par=(py, px))
    return out


def wgrad(x: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor, cin: int, cout: int, kh: int, kw: int,
          stride: int = 1, pad=(0, 0), rowscale=None, dbias=None):
    """Accumulate into dw fp32 [taps, Cout, Cin] (and dbias fp32 [Cout])."""
    native.wgrad_raw([(x, dy)], dw, cin, x.shape[-1], cout, dy.shape[-1], kh, kw, stride, pad[0], pad[1],
                     rowscale=rowscale, dbias=dbias)
    return dw


def wgrad_group(xs, dys, dw, cin, cout, kh, kw, pad=(1, 1), rowscale=None, dbias=None):
    native.wgrad_raw(list(zip(xs, dys)), dw, cin, xs[0].shape[-1], cout, dys[0].shape[-1], kh, kw, 1, pad[0], pad[1],
                     rowscale=rowscale, dbias=dbias)
    return dw
