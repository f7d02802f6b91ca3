"""Backbone conv shapes for ONE `ncu --set full` capture (round-2 investigation of the layers that sit far from both
rooflines): res2 3x3 64->64 (+shift+ReLU), res2 1x1 256->64, res2 1x1 64->256 + residual, res2 dgrad 1x1 64->256 with
residual + mask, res3 3x3 128->128, all at batch 8 of 800x1333."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.ops import native

dev, bf = "cuda", torch.bfloat16
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
t64 = torch.randn(8, 200, 334, 64, device=dev).to(bf)
o64 = torch.empty_like(t64)
t256 = torch.randn(8, 200, 334, 256, device=dev).to(bf)
o256 = torch.empty_like(t256)
m256 = torch.randn(8, 200, 334, 256, device=dev).to(bf)
w33 = (torch.randn(9, 64, 64, device=dev) * 0.05).to(bf)
w_a = (torch.randn(1, 64, 256, device=dev) * 0.05).to(bf)
w_c = (torch.randn(1, 256, 64, device=dev) * 0.05).to(bf)
b = torch.zeros(256, device=dev)
t128 = torch.randn(8, 100, 167, 128, device=dev).to(bf)
o128 = torch.empty_like(t128)
w128 = (torch.randn(9, 128, 128, device=dev) * 0.03).to(bf)
for _ in range(reps):
    native.conv_raw(0, [(t64, o64, None, None)], w33, 64, 64, 64, 64, 3, 3, 1, 1, 1, shift=b[:64].contiguous(), relu=True)
    native.conv_raw(0, [(t256, o64, None, None)], w_a, 256, 256, 64, 64, 1, 1, 1, 0, 0, shift=b[:64].contiguous(), relu=True)
    native.conv_raw(0, [(t64, o256, t256, None)], w_c, 64, 64, 256, 256, 1, 1, 1, 0, 0, shift=b, relu=True)
    native.conv_raw(1, [(t64, o256, t256, m256)], w_c, 64, 64, 256, 256, 1, 1, 1, 0, 0)
    native.conv_raw(0, [(t128, o128, None, None)], w128, 128, 128, 128, 128, 3, 3, 1, 1, 1, shift=b[:128].contiguous(), relu=True)
torch.cuda.synchronize()
print("done")
