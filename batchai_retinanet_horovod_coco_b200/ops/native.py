"""ctypes bindings for ``libb200ops.so`` (the hand-written sm_100a kernels, ``csrc/*.cu``).

The library is built in-tree by ``build_native.py`` (``__graft_entry__.build()``) and loaded
from ``batchai_retinanet_horovod_coco_b200/_native/``.  On a machine with a CUDA device the ops
**fail loudly** if the library is missing (no silent PyTorch fallback); on CPU-only machines
``available()`` is False and callers use the plain-PyTorch spec implementations.

All wrappers take ``torch`` CUDA tensors, pass raw pointers + the current stream, and never
synchronise.  Layout convention: activations are NHWC bf16 tensors of shape ``[N, H, W, C]``.
"""

from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_LIB = None
_LOCK = threading.Lock()
_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_native", "libb200ops.so")

MAX_PROBLEMS = 5


class ConvProblem(C.Structure):
    _fields_ = [("inp", C.c_void_p), ("out", C.c_void_p), ("add", C.c_void_p), ("mask", C.c_void_p),
                ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("OHo", C.c_int), ("OWo", C.c_int),
                ("colsum", C.c_void_p)]


class ConvArgs(C.Structure):
    _fields_ = [("mode", C.c_int), ("num_problems", C.c_int), ("prob", ConvProblem * MAX_PROBLEMS),
                ("weight", C.c_void_p), ("K", C.c_int), ("ldi", C.c_int), ("Nrows", C.c_int), ("ldw", C.c_int),
                ("ldo", C.c_int), ("KH", C.c_int), ("KW", C.c_int), ("stride", C.c_int), ("pad_h", C.c_int),
                ("pad_w", C.c_int), ("par_y", C.c_int), ("par_x", C.c_int), ("shift", C.c_void_p),
                ("relu", C.c_int), ("out_fp32", C.c_int), ("add_after_mask", C.c_int)]


class WgradProblem(C.Structure):
    _fields_ = [("x", C.c_void_p), ("dy", C.c_void_p), ("N", C.c_int), ("H", C.c_int), ("W", C.c_int),
                ("OH", C.c_int), ("OW", C.c_int)]


class WgradArgs(C.Structure):
    _fields_ = [("num_problems", C.c_int), ("prob", WgradProblem * MAX_PROBLEMS), ("dw", C.c_void_p),
                ("Cin", C.c_int), ("ldx", C.c_int), ("Cout", C.c_int), ("ldy", C.c_int), ("KH", C.c_int),
                ("KW", C.c_int), ("stride", C.c_int), ("pad_h", C.c_int), ("pad_w", C.c_int),
                ("rowscale", C.c_void_p), ("dbias", C.c_void_p)]


class WeightEntry(C.Structure):
    _fields_ = [("src", C.c_void_p), ("scale", C.c_void_p), ("dst_f", C.c_void_p), ("dst_d", C.c_void_p),
                ("taps", C.c_int), ("Cout", C.c_int), ("Cin", C.c_int), ("ldf", C.c_int), ("ldd", C.c_int),
                ("block_begin", C.c_longlong)]


class ArHostArgs(C.Structure):
    _fields_ = [("bufs", C.c_void_p * 8), ("flags", C.c_void_p * 8), ("prescale", C.c_void_p * 8),
                ("outs", C.c_void_p * 8), ("out", C.c_void_p), ("rank", C.c_int), ("world", C.c_int), ("offset", C.c_longlong),
                ("count", C.c_longlong), ("postscale", C.c_float), ("epoch", C.c_uint32), ("algo", C.c_int),
                ("max_blocks", C.c_int), ("has_prescale", C.c_int), ("epoch_ptr", C.c_void_p), ("mc_buf", C.c_void_p), ("mc_out", C.c_void_p),
                ("external_sync", C.c_int), ("stage", C.c_void_p * 8), ("wire_bf16", C.c_int),
                ("err", C.c_void_p), ("timeout_cycles", C.c_longlong)]


def lib_path() -> str:
    return _LIB_PATH


def _load():
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(f"native library missing: {_LIB_PATH}; run `python build_native.py`")
        lib = C.CDLL(_LIB_PATH)
        lib.b200_last_error.restype = C.c_char_p
        lib.b200_launch_count.restype = C.c_longlong
        lib.b200_weight_table_blocks.restype = C.c_longlong
        lib.b200_filter_detections_workspace.restype = C.c_longlong
        lib.b200_flag_pad_bytes.restype = C.c_longlong
        _LIB = lib
        return lib


def available() -> bool:
    """True when CUDA kernels can run here (a CUDA device is visible)."""
    return torch.cuda.is_available()


def require():
    lib = _load()
    if not torch.cuda.is_available():
        raise RuntimeError("native ops need a CUDA device")
    return lib


def _check(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {_load().b200_last_error().decode()}")


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def deterministic() -> bool:
    """B200_DETERMINISTIC=1 at load time: bit-reproducible (slow) accumulation order in every kernel."""
    return bool(_load().b200_deterministic())


def launch_count() -> int:
    return int(_load().b200_launch_count())


def add_launch_count(n: int):
    """Report kernels executed through a CUDA-graph replay (they bypass the launch wrappers)."""
    _load().b200_add_launch_count(C.c_longlong(int(n)))


def reset_launch_count():
    _load().b200_reset_launch_count()


# ------------------------------------------------------------------------------ convolution
def same_pad_before(n: int, k: int, s: int) -> int:
    out = -(-n // s)
    total = max((out - 1) * s + k - n, 0)
    return total // 2


def conv_raw(mode: int, problems, weight: torch.Tensor, K: int, ldi: int, Nrows: int, ldo: int, KH: int, KW: int,
             stride: int, pad_h: int, pad_w: int, shift=None, relu=False, out_fp32=False, par=(0, 0),
             add_after_mask=False):
    """problems: list of (inp, out, add, mask[, colsum]) NHWC tensors (colsum: fp32 [Nrows] vector that receives
    += the sum of ``out`` over pixels).  weight: bf16 [taps, Nrows, ldw]."""
    lib = require()
    a = ConvArgs()
    a.mode = mode
    a.num_problems = len(problems)
    for i, prob in enumerate(problems):
        inp, out, add, mask = prob[:4]
        p = a.prob[i]
        p.colsum = 0 if len(prob) < 5 or prob[4] is None else prob[4].data_ptr()
        p.inp, p.out = inp.data_ptr(), out.data_ptr()
        p.add = 0 if add is None else add.data_ptr()
        p.mask = 0 if mask is None else mask.data_ptr()
        p.N, p.H, p.W = inp.shape[0], inp.shape[1], inp.shape[2]
        p.OHo, p.OWo = out.shape[1], out.shape[2]
    a.weight = weight.data_ptr()
    a.K, a.ldi, a.Nrows, a.ldw, a.ldo = K, ldi, Nrows, weight.shape[-1], ldo
    a.KH, a.KW, a.stride, a.pad_h, a.pad_w = KH, KW, stride, pad_h, pad_w
    a.par_y, a.par_x = par
    a.shift = 0 if shift is None else shift.data_ptr()
    a.relu = int(relu)
    a.out_fp32 = int(out_fp32)
    a.add_after_mask = int(add_after_mask)
    _check(lib.b200_conv(C.byref(a), _stream()), "b200_conv")


def conv_set_trace(buf):
    """buf: int64 device tensor [>= SMs * 16] receiving the conv kernel's in-kernel timeline, or None (off)."""
    lib = require()
    lib.b200_conv_set_trace.restype = None
    lib.b200_conv_set_trace.argtypes = [C.c_void_p]
    lib.b200_conv_set_trace(C.c_void_p(0 if buf is None else buf.data_ptr()))


def wgrad_raw(problems, dw: torch.Tensor, Cin: int, ldx: int, Cout: int, ldy: int, KH: int, KW: int, stride: int,
              pad_h: int, pad_w: int, rowscale=None, dbias=None):
    """problems: list of (x NHWC, dy NHWC).  dw fp32 [taps, Cout, Cin] accumulated into."""
    lib = require()
    a = WgradArgs()
    a.num_problems = len(problems)
    for i, (x, dy) in enumerate(problems):
        p = a.prob[i]
        p.x, p.dy = x.data_ptr(), dy.data_ptr()
        p.N, p.H, p.W = x.shape[0], x.shape[1], x.shape[2]
        p.OH, p.OW = dy.shape[1], dy.shape[2]
    a.dw = dw.data_ptr()
    a.Cin, a.ldx, a.Cout, a.ldy = Cin, ldx, Cout, ldy
    a.KH, a.KW, a.stride, a.pad_h, a.pad_w = KH, KW, stride, pad_h, pad_w
    a.rowscale = 0 if rowscale is None else rowscale.data_ptr()
    a.dbias = 0 if dbias is None else dbias.data_ptr()
    _check(lib.b200_wgrad(C.byref(a), _stream()), "b200_wgrad")


def stem_conv(x4, w_packed, shift, y, relu=True):
    """x4 bf16 [N,H,W,4] -> y bf16 [N,ceil(H/2),ceil(W/2),64] = relu(conv7x7/2(x) + shift)  (relu optional)."""
    n, h, w, c = x4.shape
    assert c == 4 and w_packed.shape == (64, 256)
    _check(require().b200_stem_conv_ex(_ptr(x4), _ptr(w_packed), _ptr(shift), _ptr(y), n, h, w, int(relu), _stream()),
           "b200_stem_conv")


def stem_wgrad(x4, dy, scale, dw):
    """dw fp32 [49,64,3] += scale[co] * grad of the stem filter (x4 bf16 [N,H,W,4], dy bf16 [N,H/2,W/2,64])."""
    n, h, w, c = x4.shape
    _check(require().b200_stem_wgrad(_ptr(x4), _ptr(dy), _ptr(scale), _ptr(dw), n, h, w, _stream()), "b200_stem_wgrad")


def pack_stem_weight(master_tci, scale=None):
    """master fp32 [49, 64, 3] (tap, co, ci) -> bf16 [64, 256], k = r*28 + s*4 + c, BN scale folded."""
    w = master_tci.view(7, 7, 64, 3).permute(2, 0, 1, 3)          # co, r, s, c
    if scale is not None:
        w = w * scale.view(64, 1, 1, 1)
    out = torch.zeros(64, 7, 7, 4, device=master_tci.device, dtype=torch.float32)
    out[..., :3] = w
    full = torch.zeros(64, 256, device=master_tci.device, dtype=torch.bfloat16)
    full[:, :196] = out.reshape(64, 196).to(torch.bfloat16)
    return full


class WeightTable:
    """Device table driving the one-launch master->bf16 working-weight preparation."""

    def __init__(self, entries):
        lib = require()
        n = len(entries)
        self.n = n
        self.host = (WeightEntry * n)()
        for i, e in enumerate(entries):
            h = self.host[i]
            h.src, h.scale = e["src"].data_ptr(), (0 if e.get("scale") is None else e["scale"].data_ptr())
            h.dst_f = 0 if e.get("dst_f") is None else e["dst_f"].data_ptr()
            h.dst_d = 0 if e.get("dst_d") is None else e["dst_d"].data_ptr()
            h.taps, h.Cout, h.Cin = e["taps"], e["Cout"], e["Cin"]
            h.ldf = 0 if e.get("dst_f") is None else e["dst_f"].shape[-1]
            h.ldd = 0 if e.get("dst_d") is None else e["dst_d"].shape[-1]
        self.blocks = int(lib.b200_weight_table_blocks(self.host, n))
        raw = bytes(self.host)
        self.dev = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
        self._keep = entries

    def run(self):
        _check(_load().b200_prepare_weights(self.host, self.n, _ptr(self.dev), _stream()), "b200_prepare_weights")


# ------------------------------------------------------------------------------ elementwise
def bn_relu_maxpool_fwd(x, scale, shift, y, idx, pad_h, pad_w):
    n, h, w, c = x.shape
    _check(_load().b200_bn_relu_maxpool_fwd(_ptr(x), _ptr(scale), _ptr(shift), _ptr(y), _ptr(idx), n, h, w, c,
                                            y.shape[1], y.shape[2], pad_h, pad_w, _stream()), "bn_relu_maxpool_fwd")


def bn_relu_maxpool_bwd(idx, scale, dy, dx, pad_h, pad_w):
    n, h, w, c = dx.shape
    _check(_load().b200_bn_relu_maxpool_bwd(_ptr(idx), _ptr(scale), _ptr(dy), _ptr(dx), n, h, w, c, dy.shape[1],
                                            dy.shape[2], pad_h, pad_w, _stream()), "bn_relu_maxpool_bwd")


def upsample_add(lateral, top, out):
    n, h, w, c = lateral.shape
    _check(_load().b200_upsample_add(_ptr(lateral), _ptr(top), _ptr(out), n, h, w, c, top.shape[1], top.shape[2],
                                     _stream()), "upsample_add")


def upsample_bwd_add(dmerged, dtop):
    n, h, w, c = dmerged.shape
    _check(_load().b200_upsample_bwd_add(_ptr(dmerged), _ptr(dtop), n, h, w, c, dtop.shape[1], dtop.shape[2],
                                         _stream()), "upsample_bwd_add")


def relu(x, y):
    _check(_load().b200_relu(_ptr(x), _ptr(y), C.c_longlong(x.numel()), _stream()), "relu")


def relu_bwd(y, dy, dx):
    _check(_load().b200_relu_bwd(_ptr(y), _ptr(dy), _ptr(dx), C.c_longlong(y.numel()), _stream()), "relu_bwd")


def add_mask(a, b, mask, out):
    _check(_load().b200_add_mask(_ptr(a), _ptr(b), _ptr(mask), _ptr(out), C.c_longlong(a.numel()), _stream()), "add_mask")


def colsum(x2d, cols, out_accum):
    rows = x2d.numel() // x2d.shape[-1]
    _check(_load().b200_colsum(_ptr(x2d), C.c_longlong(rows), cols, x2d.shape[-1], _ptr(out_accum), _stream()), "colsum")


def preprocess(src_u8, dst, mean, img_hw=None, scale: float = 1.0):
    """uint8 [N,H,W,C] -> bf16 [N,H,W,ld] minus mean; pixels outside each image's (h, w) (int32 [N,2], optional)
    become exactly 0, like the reference's zero padding after normalisation."""
    n, h, w, c = src_u8.shape
    m = (C.c_float * 4)(*([float(v) for v in mean] + [0.0] * (4 - len(mean))))
    _check(_load().b200_preprocess_scaled(_ptr(src_u8), _ptr(dst), n, h, w, c, dst.shape[-1], m, _ptr(img_hw),
                                          C.c_float(scale), _stream()), "preprocess")


# ------------------------------------------------------------------------------ targets / losses
def generate_anchors(out, H, W, stride, size):
    _check(_load().b200_generate_anchors(_ptr(out), H, W, stride, C.c_float(size), _stream()), "generate_anchors")


# ------------------------------------------------------------------------------ MobileNet / DenseNet layers
def dwconv3x3(mode: int, a, b, out, in_hw, out_hw, stride: int, pad):
    """mode 0: y = dw(x, w); 1: dx = dw^T(dy, w); 2: dw += x (*) dy.  w / dw fp32 [C][9]; in_hw = (H, W) of x."""
    n, c = (a.shape[0], a.shape[-1])
    _check(_load().b200_dwconv3x3(mode, _ptr(a), _ptr(b), _ptr(out), n, in_hw[0], in_hw[1], c, out_hw[0], out_hw[1], stride,
                                  pad[0], pad[1], _stream()), "b200_dwconv3x3")


def bn_stats(x, channels, sums, mean, var, moving_mean=None, moving_var=None, momentum=0.99):
    """x [..., C] may be a channel slice of a wider buffer: the channel stride is x.stride(-2)."""
    rows = x.numel() // x.shape[-1]
    import ctypes
    _check(_load().b200_bn_stats(_ptr(x), ctypes.c_longlong(rows), int(channels), x.stride(-2), _ptr(sums), _ptr(mean), _ptr(var),
                                 _ptr(moving_mean), _ptr(moving_var), ctypes.c_float(momentum), _stream()), "b200_bn_stats")


def bn_act_fwd(x, mean, var, gamma, beta, eps, y, channels, act):
    rows = x.numel() // x.shape[-1]
    _check(_load().b200_bn_act_fwd(_ptr(x), _ptr(mean), _ptr(var), _ptr(gamma), _ptr(beta), C.c_float(eps), _ptr(y),
                                   C.c_longlong(rows), channels, x.stride(-2), y.stride(-2), act, _stream()), "b200_bn_act_fwd")


def bn_act_bwd(x, dy, mean, var, gamma, beta, eps, sums, channels, act, dx, dgamma, dbeta, accumulate=False):
    rows = x.numel() // x.shape[-1]
    _check(_load().b200_bn_act_bwd(_ptr(x), _ptr(dy), _ptr(mean), _ptr(var), _ptr(gamma), _ptr(beta), C.c_float(eps), _ptr(sums),
                                   C.c_longlong(rows), channels, x.stride(-2), dy.stride(-2), dx.stride(-2), act, int(accumulate),
                                   _ptr(dx), _ptr(dgamma), _ptr(dbeta), _stream()), "b200_bn_act_bwd")


def avgpool2(x, y, backward=False, channels=None):
    """forward: x [N,H,W,C] -> y [N,H/2,W/2,C]; backward: x = dy [N,H/2,W/2,C] -> y = dx [N,H,W,C].  Either tensor may
    be a channel slice of a wider buffer (its stride(2) is the channel stride)."""
    big = y if backward else x
    n, h, w, c = big.shape
    c = channels or min(x.shape[-1], y.shape[-1])
    _check(_load().b200_avgpool2(int(backward), _ptr(x), _ptr(y), n, h, w, c, x.stride(2), y.stride(2), _stream()),
           "b200_avgpool2")


def wgrad_small(x, dy, dw, cin, cout, kh, kw, stride, pad):
    n, h, w, ldx = x.shape
    _check(_load().b200_wgrad_small(_ptr(x), _ptr(dy), _ptr(dw), n, h, w, cin, ldx, dy.shape[1], dy.shape[2], cout,
                                    dy.shape[-1], kh, kw, stride, pad[0], pad[1], _stream()), "b200_wgrad_small")


def resize_linear_u8(src, dst, scale_x: float, scale_y: float):
    """src uint8 [sh, sw, 3] (device, row pitch = stride(0)) -> dst uint8 view [dh, dw, 3] (e.g. a slot of the batch
    tensor).  scale_* = source pixels per destination pixel (1 / fx of cv2.resize)."""
    _check(_load().b200_resize_linear_u8(_ptr(src), src.shape[0], src.shape[1], C.c_longlong(src.stride(0)), _ptr(dst),
                                         dst.shape[0], dst.shape[1], C.c_longlong(dst.stride(0)), C.c_double(scale_x),
                                         C.c_double(scale_y), _stream()), "resize_linear_u8")


def invert_affine_like_cv2(m2x3):
    """The dst -> src map cv2.warpAffine derives from the forward matrix it is given (same float64 operations)."""
    import numpy as np
    M = np.asarray(m2x3, dtype=np.float64).reshape(2, 3).copy()
    D = M[0, 0] * M[1, 1] - M[0, 1] * M[1, 0]
    D = 1.0 / D if D != 0 else 0.0
    A11, A22 = M[1, 1] * D, M[0, 0] * D
    M[0, 0] = A11
    M[0, 1] *= -D
    M[1, 0] *= -D
    M[1, 1] = A22
    b1 = -M[0, 0] * M[0, 2] - M[0, 1] * M[1, 2]
    b2 = -M[1, 0] * M[0, 2] - M[1, 1] * M[1, 2]
    M[0, 2], M[1, 2] = b1, b2
    return M


def warp_affine_u8(src, dst, forward_2x3, border_replicate: bool = True, cval: int = 0):
    """cv2.warpAffine(src, forward_2x3, (dw, dh), INTER_LINEAR, BORDER_REPLICATE | BORDER_CONSTANT) on the device."""
    inv = invert_affine_like_cv2(forward_2x3).reshape(-1)
    arr = (C.c_double * 6)(*[float(v) for v in inv])
    _check(_load().b200_warp_affine_u8(_ptr(src), src.shape[0], src.shape[1], C.c_longlong(src.stride(0)), _ptr(dst),
                                       dst.shape[0], dst.shape[1], C.c_longlong(dst.stride(0)), arr,
                                       int(border_replicate), int(cval), _stream()), "warp_affine_u8")


def anchors_for_shape_gpu(feature_shapes, device) -> torch.Tensor:
    from .anchors import SIZES, STRIDES
    total = sum(h * w * 9 for h, w in feature_shapes)
    out = torch.empty(total, 4, device=device, dtype=torch.float32)
    off = 0
    for (h, w), s, z in zip(feature_shapes, STRIDES, SIZES):
        generate_anchors(out[off:], h, w, s, z)
        off += h * w * 9
    return out


def assign_targets(anchors, gt, gt_count, img_hw, state, cls, reg_target, num_pos, neg_thr=0.4, pos_thr=0.5):
    A = anchors.shape[0]
    B, Gmax = gt.shape[0], gt.shape[1]
    _check(_load().b200_assign_targets(_ptr(anchors), A, _ptr(gt), _ptr(gt_count), _ptr(img_hw), B, Gmax,
                                       C.c_float(neg_thr), C.c_float(pos_thr), _ptr(state), _ptr(cls),
                                       _ptr(reg_target), _ptr(num_pos), _stream()), "assign_targets")


def focal_loss(logits, dlogits, K, state, cls, A_total, a_off, num_pos, loss_accum, alpha=0.25, gamma=2.0,
               loss_scale=1.0):
    """logits NHWC [B,H,W,ld] of one level (ld >= 9K)."""
    B, H, W, ld = logits.shape
    _check(_load().b200_focal_loss(_ptr(logits), _ptr(dlogits), B, H * W * 9, K, ld, _ptr(state), _ptr(cls), A_total,
                                   a_off, _ptr(num_pos), C.c_float(alpha), C.c_float(gamma), C.c_float(loss_scale),
                                   _ptr(loss_accum), _stream()), "focal_loss")


def smooth_l1(pred, dpred, state, reg_target, A_total, a_off, num_pos, loss_accum, sigma=3.0, loss_scale=1.0):
    B, H, W, ld = pred.shape
    _check(_load().b200_smooth_l1(_ptr(pred), _ptr(dpred), B, H * W * 9, ld, _ptr(state), _ptr(reg_target), A_total,
                                  a_off, _ptr(num_pos), C.c_float(sigma), C.c_float(loss_scale), _ptr(loss_accum),
                                  _stream()), "smooth_l1")


# ------------------------------------------------------------------------------ optimizer
def sumsq(x, out):
    _check(_load().b200_sumsq(_ptr(x), C.c_longlong(x.numel()), _ptr(out), _stream()), "sumsq")


def clip_scale(sumsq_t, clipnorm, scale):
    _check(_load().b200_clip_scale(_ptr(sumsq_t), C.c_float(clipnorm), _ptr(scale), _stream()), "clip_scale")


def maxpool2_fwd(x, y):
    """2x2/2 max-pool, valid padding: x bf16 [N,H,W,C] -> y [N,H//2,W//2,C]."""
    n, h, w, c = x.shape
    _check(require().b200_maxpool2_fwd(_ptr(x), _ptr(y), n, h, w, c, _stream()), "maxpool2_fwd")


def maxpool2_relu_bwd(x, dy, dx):
    """dx = gradient w.r.t. the pre-activation of the ReLU output ``x`` that was pooled (dy: gradient of the pool)."""
    n, h, w, c = x.shape
    _check(require().b200_maxpool2_relu_bwd(_ptr(x), _ptr(dy), _ptr(dx), n, h, w, c, _stream()), "maxpool2_relu_bwd")


def adam(p, g, m, v, grad_scale, lr_t, beta1, beta2, eps):
    """``lr_t`` may be a float or a 1-element fp32 device tensor (read on the device: CUDA-graph friendly)."""
    if isinstance(lr_t, torch.Tensor):
        _check(_load().b200_adam_devlr(_ptr(p), _ptr(g), _ptr(m), _ptr(v), C.c_longlong(p.numel()), _ptr(grad_scale),
                                       _ptr(lr_t), C.c_float(beta1), C.c_float(beta2), C.c_float(eps), _stream()), "adam")
        return
    _check(_load().b200_adam(_ptr(p), _ptr(g), _ptr(m), _ptr(v), C.c_longlong(p.numel()), _ptr(grad_scale),
                             C.c_float(lr_t), C.c_float(beta1), C.c_float(beta2), C.c_float(eps), _stream()), "adam")


# ------------------------------------------------------------------------------ detection tail
_DET_WS = {}


def filter_detections_levels(cls_levels, reg_levels, K, anchors, img_hw, score_thr=0.05, nms_thr=0.5, max_det=300,
                             max_cand=65536):
    """cls_levels / reg_levels: lists of NHWC bf16 tensors (logits / offsets) per pyramid level."""
    lib = require()
    B = cls_levels[0].shape[0]
    dev = cls_levels[0].device
    nl = len(cls_levels)
    key = (B, max_cand, dev.index)
    if key not in _DET_WS:
        nbytes = int(lib.b200_filter_detections_workspace(B, max_cand))
        _DET_WS[key] = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    ws = _DET_WS[key]
    cl = (C.c_void_p * nl)(*[t.data_ptr() for t in cls_levels])
    rl = (C.c_void_p * nl)(*[t.data_ptr() for t in reg_levels])
    la = (C.c_int * nl)(*[t.shape[1] * t.shape[2] * 9 for t in cls_levels])
    out_b = torch.empty(B, max_det, 4, device=dev, dtype=torch.float32)
    out_s = torch.empty(B, max_det, device=dev, dtype=torch.float32)
    out_l = torch.empty(B, max_det, device=dev, dtype=torch.int32)
    _check(lib.b200_filter_detections(cl, rl, la, nl, cls_levels[0].shape[-1], reg_levels[0].shape[-1], K,
                                      _ptr(anchors), anchors.shape[0], B, int(img_hw[0]), int(img_hw[1]),
                                      C.c_float(score_thr), C.c_float(nms_thr), max_det, max_cand, _ptr(ws),
                                      C.c_longlong(ws.numel()), _ptr(out_b), _ptr(out_s), _ptr(out_l), _stream()),
           "filter_detections")
    return out_b, out_s, out_l.long()


def filter_detections(boxes, classification, score_thr=0.05, max_det=300, nms_thr=0.5):
    """Generic-tensor entry used by the spec-layer prediction graph on CUDA: boxes (B,A,4) fp32 already
    decoded, classification (B,A,K) probabilities.  Threshold/sort in torch, NMS by the native kernel."""
    B = boxes.shape[0]
    outs = []
    for b in range(B):
        cand = (classification[b] > score_thr).nonzero(as_tuple=False)
        a_idx, c_idx = cand[:, 0], cand[:, 1]
        sc = classification[b][a_idx, c_idx]
        keep = nms(boxes[b][a_idx], sc, nms_thr, max_det, c_idx)
        a_idx, c_idx, sc = a_idx[keep], c_idx[keep], sc[keep]
        k = min(max_det, sc.numel())
        ob = torch.full((max_det, 4), -1.0, device=boxes.device, dtype=boxes.dtype)
        os_ = torch.full((max_det,), -1.0, device=boxes.device, dtype=classification.dtype)
        ol = torch.full((max_det,), -1, device=boxes.device, dtype=torch.int64)
        ob[:k], os_[:k], ol[:k] = boxes[b][a_idx[:k]], sc[:k], c_idx[:k]
        outs.append((ob, os_, ol))
    return (torch.stack([o[0] for o in outs]), torch.stack([o[1] for o in outs]), torch.stack([o[2] for o in outs]))


def nms(boxes, scores, thr, max_output=None, classes=None):
    """Greedy NMS (optionally class-aware).  Returns kept indices in descending-score order."""
    lib = require()
    n = boxes.shape[0]
    if n == 0:
        return torch.zeros(0, dtype=torch.int64, device=boxes.device)
    order = torch.argsort(scores, descending=True, stable=True)
    b = boxes[order].contiguous().float()
    cl = None if classes is None else classes[order].contiguous().int()
    words = (n + 63) // 64
    mask = torch.empty(n * words, dtype=torch.int64, device=boxes.device)
    keep = torch.zeros(n, dtype=torch.int32, device=boxes.device)
    _check(lib.b200_nms_sorted(_ptr(b), _ptr(cl), n, C.c_float(thr), _ptr(mask), _ptr(keep), _stream()), "nms_sorted")
    idx = order[keep.bool()]
    if max_output is not None:
        idx = idx[:max_output]
    return idx


# ------------------------------------------------------------------------------ collectives
def flag_pad_bytes() -> int:
    return int(_load().b200_flag_pad_bytes())


def allreduce_host(args: ArHostArgs):
    _check(_load().b200_allreduce_host(C.byref(args), _stream()), "b200_allreduce_host")


def ar_scatter(args: ArHostArgs, lo: int, hi: int, stream=None):
    st = _stream() if stream is None else C.c_void_p(stream.cuda_stream)
    _check(_load().b200_ar_scatter(C.byref(args), C.c_longlong(lo), C.c_longlong(hi), st), "b200_ar_scatter")


def ar_reduce_staged(args: ArHostArgs):
    _check(_load().b200_ar_reduce_staged(C.byref(args), _stream()), "b200_ar_reduce_staged")


def broadcast_host(args: ArHostArgs, root: int):
    _check(_load().b200_broadcast_host(C.byref(args), root, _stream()), "b200_broadcast_host")


def ipc_alloc(nbytes: int) -> int:
    p = C.c_void_p()
    _check(_load().b200_ipc_alloc(C.byref(p), C.c_longlong(nbytes)), "ipc_alloc")
    return int(p.value)


def ipc_free(ptr: int):
    _check(_load().b200_ipc_free(C.c_void_p(ptr)), "ipc_free")


def ipc_get_handle(ptr: int) -> bytes:
    buf = C.create_string_buffer(64)
    _check(_load().b200_ipc_get_handle(C.c_void_p(ptr), buf), "ipc_get_handle")
    return bytes(buf.raw)


def ipc_open_handle(handle: bytes) -> int:
    p = C.c_void_p()
    _check(_load().b200_ipc_open_handle(C.c_char_p(handle), C.byref(p)), "ipc_open_handle")
    return int(p.value)


def ipc_close_handle(ptr: int):
    _check(_load().b200_ipc_close_handle(C.c_void_p(ptr)), "ipc_close_handle")


# ---- VMM regions + NVLink multicast (csrc/vmm_multicast.cu) ----------------------------------------
def vmm_supported(device: int):
    """(vmm_with_posix_fd, multicast) capability of ``device``."""
    mc = C.c_int(0)
    ok = _load().b200_vmm_supported(int(device), C.byref(mc))
    return bool(ok), bool(mc.value)


def vmm_granularity(device: int, world: int) -> int:
    g = C.c_longlong(0)
    _check(_load().b200_vmm_granularity(int(device), int(world), C.byref(g)), "vmm_granularity")
    return int(g.value)


def vmm_create(device: int, nbytes: int):
    """-> (handle, fd): physical allocation on ``device`` and a POSIX fd peers can import."""
    h, fd = C.c_ulonglong(0), C.c_int(-1)
    _check(_load().b200_vmm_create(int(device), C.c_longlong(nbytes), C.byref(h), C.byref(fd)), "vmm_create")
    return int(h.value), int(fd.value)


def vmm_import(fd: int) -> int:
    h = C.c_ulonglong(0)
    _check(_load().b200_vmm_import(int(fd), C.byref(h)), "vmm_import")
    return int(h.value)


def vmm_map(handle: int, nbytes: int, device: int, alignment: int) -> int:
    p = C.c_void_p()
    _check(_load().b200_vmm_map(C.c_ulonglong(handle), C.c_longlong(nbytes), int(device), C.c_longlong(alignment),
                                C.byref(p)), "vmm_map")
    return int(p.value)


def vmm_unmap(ptr: int, nbytes: int):
    _check(_load().b200_vmm_unmap(C.c_void_p(ptr), C.c_longlong(nbytes)), "vmm_unmap")


def vmm_release(handle: int):
    _check(_load().b200_vmm_release(C.c_ulonglong(handle)), "vmm_release")


def mc_create(world: int, nbytes: int):
    h, fd = C.c_ulonglong(0), C.c_int(-1)
    _check(_load().b200_mc_create(int(world), C.c_longlong(nbytes), C.byref(h), C.byref(fd)), "mc_create")
    return int(h.value), int(fd.value)


def mc_add_device(mc: int, device: int):
    _check(_load().b200_mc_add_device(C.c_ulonglong(mc), int(device)), "mc_add_device")


def mc_bind(mc: int, mem: int, nbytes: int):
    _check(_load().b200_mc_bind(C.c_ulonglong(mc), C.c_ulonglong(mem), C.c_longlong(nbytes)), "mc_bind")


def enable_peer_access(device: int, world_local: int):
    """cudaDeviceEnablePeerAccess from ``device`` (current) to devices 0..world_local-1."""
    _check(_load().b200_enable_peer_access(int(device), int(world_local)), "enable_peer_access")


class RawCudaBuffer:
    """A cudaMalloc'ed region exposed to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.ptr, self.nbytes, self.owner = ptr, nbytes, owner
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2,
                                         "strides": None}

    def tensor(self, device) -> torch.Tensor:
        t = torch.as_tensor(self, device=device)
        t._b200_owner = self
        return t
