"""Device time of the MobileNet / DenseNet layer kernels and of the small-Cin filter gradient at flagship shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from batchai_retinanet_horovod_coco_b200.ops import native
dev, bf = "cuda", torch.bfloat16
def t(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
B = 8
x8 = torch.randn(B, 800, 1333, 8, device=dev).to(bf)
for cout, stride in ((64, 1), (32, 2)):
    oh, ow = -(-800 // stride), -(-1333 // stride)
    dy = torch.randn(B, oh, ow, cout, device=dev).to(bf)
    dw = torch.zeros(9, cout, 3, device=dev)
    print(f"wgrad_small cout={cout} stride={stride}: {t(lambda: native.wgrad_small(x8, dy, dw, 3, cout, 3, 3, stride, (1, 1) if stride == 1 else (0, 0))):.3f} ms")
for (h, w, c, s) in ((400, 667, 32, 1), (400, 667, 64, 2), (200, 334, 128, 1), (100, 167, 256, 1), (50, 84, 512, 1)):
    oh, ow = -(-h // s), -(-w // s)
    x = torch.randn(B, h, w, c, device=dev).to(bf); y = torch.empty(B, oh, ow, c, device=dev, dtype=bf)
    wt = torch.randn(c, 9, device=dev); dwt = torch.zeros(c, 9, device=dev)
    mb = x.numel() * 2 / 1e6
    a = t(lambda: native.dwconv3x3(0, x, wt, y, (h, w), (oh, ow), s, (1, 1)))
    b_ = t(lambda: native.dwconv3x3(1, y, wt, x, (h, w), (oh, ow), s, (1, 1)))
    c_ = t(lambda: native.dwconv3x3(2, x, y, dwt, (h, w), (oh, ow), s, (1, 1)))
    sums = torch.zeros(2 * c, device=dev); mean = torch.zeros(c, device=dev); var = torch.ones(c, device=dev); g = torch.ones(c, device=dev)
    d_ = t(lambda: native.bn_stats(x, c, sums, mean, var))
    e_ = t(lambda: native.bn_act_fwd(x, mean, var, g, mean, 1e-3, x, c, 2))
    f_ = t(lambda: native.bn_act_bwd(x, x, mean, var, g, mean, 1e-3, sums, c, 2, x, g.clone(), g.clone()))
    print(f"{h}x{w}x{c} s{s} ({mb:.0f} MB): dw fwd {a:.3f} dgrad {b_:.3f} wgrad {c_:.3f} | bn stats {d_:.3f} fwd {e_:.3f} bwd {f_:.3f} ms")
