"""NumPy restatement of the GPU image-geometry kernels (``csrc/image_sm100.cu``): the same fixed-point arithmetic, integer
for integer.  It is the CPU oracle of the kernel math -- ``tests/test_spec_layer.py`` runs it against OpenCV over many
random sizes, scales and affine maps (which a GPU test cannot afford), ``tests/test_gpu_ops.py`` runs the kernels against
OpenCV on a few -- and documents what "matches cv2" means:

* ``resize_linear_u8``: OpenCV's 8-bit bilinear ``cv2.resize``: coefficients ``round(w * 2048)`` from a float32 fraction,
  horizontal pass in 32-bit integers, vertical pass ``(((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2``;
* ``warp_affine_u8``: ``cv2.warpAffine(..., INTER_LINEAR)``: destination -> source coordinates in 10-bit fixed point
  (``round(m * x * 1024)``), 5 fractional bits kept, weights ``32 (32 - fx)(32 - fy)`` etc. summing to 2^15, result
  ``(sum + 2^14) >> 15``; border modes ``BORDER_REPLICATE`` and ``BORDER_CONSTANT``.
"""
from __future__ import annotations

import numpy as np


def _coef(n: int, scale: float, ssize: int):
    d = np.arange(n, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int32)
    f = (f - s.astype(np.float32)).astype(np.float32)
    lo = s < 0
    s[lo], f[lo] = 0, 0.0
    hi = s >= ssize - 1
    s[hi], f[hi] = ssize - 1, 0.0
    s1 = np.minimum(s + 1, ssize - 1)
    c0 = np.rint((np.float32(1.0) - f) * np.float32(2048.0)).astype(np.int32)
    c1 = np.rint(f * np.float32(2048.0)).astype(np.int32)
    return s, s1, c0, c1


def resize_linear_u8(src: np.ndarray, dh: int, dw: int, scale_x: float, scale_y: float) -> np.ndarray:
    """src uint8 [sh, sw, 3] -> uint8 [dh, dw, 3]; scale_* = source pixels per destination pixel."""
    sh, sw = src.shape[:2]
    x0, x1, a0, a1 = _coef(dw, scale_x, sw)
    y0, y1, b0, b1 = _coef(dh, scale_y, sh)
    s = src.astype(np.int32)
    r0, r1 = s[y0], s[y1]                                     # [dh, sw, 3]
    S0 = r0[:, x0] * a0[None, :, None] + r0[:, x1] * a1[None, :, None]
    S1 = r1[:, x0] * a0[None, :, None] + r1[:, x1] * a1[None, :, None]
    v = (((b0[:, None, None] * (S0 >> 4)) >> 16) + ((b1[:, None, None] * (S1 >> 4)) >> 16) + 2) >> 2
    return np.clip(v, 0, 255).astype(np.uint8)


def warp_affine_u8(src: np.ndarray, inverse_2x3, border_replicate: bool = True, cval: int = 0) -> np.ndarray:
    """dst[y, x] = bilinear(src, inverse @ (x, y, 1)) with OpenCV's fixed-point arithmetic; same size as src."""
    sh, sw = src.shape[:2]
    m = np.asarray(inverse_2x3, dtype=np.float64).reshape(6)
    x = np.arange(sw, dtype=np.float64)
    y = np.arange(sh, dtype=np.float64)
    adelta = np.rint(m[0] * x * 1024.0).astype(np.int64)
    bdelta = np.rint(m[3] * x * 1024.0).astype(np.int64)
    X0 = np.rint((m[1] * y + m[2]) * 1024.0).astype(np.int64) + 16
    Y0 = np.rint((m[4] * y + m[5]) * 1024.0).astype(np.int64) + 16
    X = (X0[:, None] + adelta[None, :]) >> 5
    Y = (Y0[:, None] + bdelta[None, :]) >> 5
    sx, sy, fx, fy = X >> 5, Y >> 5, X & 31, Y & 31
    w00, w01 = 32 * (32 - fx) * (32 - fy), 32 * fx * (32 - fy)
    w10, w11 = 32 * (32 - fx) * fy, 32 * fx * fy
    s = src.astype(np.int64)

    def px(yy, xx):
        if border_replicate:
            return s[np.clip(yy, 0, sh - 1), np.clip(xx, 0, sw - 1)]
        ok = (yy >= 0) & (yy < sh) & (xx >= 0) & (xx < sw)
        v = s[np.clip(yy, 0, sh - 1), np.clip(xx, 0, sw - 1)]
        return np.where(ok[..., None], v, cval)

    v = (px(sy, sx) * w00[..., None] + px(sy, sx + 1) * w01[..., None] + px(sy + 1, sx) * w10[..., None] +
         px(sy + 1, sx + 1) * w11[..., None] + 16384) >> 15
    return np.clip(v, 0, 255).astype(np.uint8)
