"""The hot kernels at their flagship shapes, a few launches each, for ONE `ncu --set full` capture:
head conv fprop / dgrad(+mask) and head wgrad (3x3 256->256 @ 8x100x167), stage-2 1x1 + residual, stem conv and
stem wgrad (8x800x1333)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.ops import conv as cv, native

dev = "cuda"
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
bf = torch.bfloat16
x = torch.randn(8, 100, 167, 256, device=dev).to(bf)
w = (torch.randn(9, 256, 256, device=dev) * 0.02).to(bf)
b = torch.zeros(256, device=dev)
y = torch.empty_like(x)
dy = torch.randn(8, 100, 167, 256, device=dev).to(bf)
dw = torch.zeros(9, 256, 256, device=dev)
x2 = torch.randn(8, 200, 334, 64, device=dev).to(bf)
w2 = (torch.randn(1, 256, 64, device=dev) * 0.05).to(bf)
res = torch.randn(8, 200, 334, 256, device=dev).to(bf)
y2 = torch.empty_like(res)
x4 = torch.randn(8, 800, 1333, 4, device=dev).to(bf)
wp = native.pack_stem_weight(torch.randn(49, 64, 3, device=dev) * 0.05)
ys = torch.empty(8, 400, 667, 64, device=dev, dtype=bf)
dws = torch.zeros(49, 64, 3, device=dev)
ones = torch.ones(64, device=dev)
for _ in range(reps):
    native.conv_raw(0, [(x, y, None, None)], w, 256, 256, 256, 256, 3, 3, 1, 1, 1, shift=b, relu=True)       # head fprop
    native.conv_raw(1, [(dy, y, None, x)], w, 256, 256, 256, 256, 3, 3, 1, 1, 1)                            # head dgrad + mask
    native.wgrad_raw([(x, dy)], dw, 256, 256, 256, 256, 3, 3, 1, 1, 1, dbias=None)                          # head wgrad
    native.conv_raw(0, [(x2, y2, res, None)], w2, 64, 64, 256, 256, 1, 1, 1, 0, 0, shift=b, relu=True)      # res2 branch2c
    native.stem_conv(x4, wp, b[:64].contiguous(), ys)
    native.stem_wgrad(x4, ys, ones, dws)
torch.cuda.synchronize()
print("done")
