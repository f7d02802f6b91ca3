"""Micro-benchmark of the tcgen05 conv kernels against cuDNN (bf16 channels_last) on the
RetinaNet hot shapes.  CUDA-event timing, warm-up, L2 flushed between iterations."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from batchai_retinanet_horovod_coco_b200.ops import conv as cv

dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


CASES = [  # name, n, h, w, cin, cout, k
    ("head3x3_P3", 8, 100, 167, 256, 256, 3),
    ("head3x3_P4", 8, 50, 84, 256, 256, 3),
    ("cls_out_P3", 8, 100, 167, 256, 720, 3),
    ("res2_1x1_64_256", 8, 200, 334, 64, 256, 1),
    ("res2_3x3_64", 8, 200, 334, 64, 64, 3),
    ("res3_3x3_128", 8, 100, 167, 128, 128, 3),
    ("res4_1x1_1024_256", 8, 50, 84, 1024, 256, 1),
    ("res4_3x3_256", 8, 50, 84, 256, 256, 3),
    ("res5_1x1_512_2048", 8, 25, 42, 512, 2048, 1),
]
rows = []
for name, n, h, w, cin, cout, k in CASES:
    x = torch.randn(n, h, w, cin, device=dev).to(torch.bfloat16)
    wt = torch.randn(cout, cin, k, k, device=dev) * 0.02
    wf, wd = cv.pack_working(cv.pack_master(wt))
    out = torch.empty(n, h, w, cv.round8(cout), device=dev, dtype=torch.bfloat16)
    p = k // 2
    flops = 2.0 * n * h * w * cin * cout * k * k
    t_f = timeit(lambda: cv.fprop(x, wf, cin, cout, k, k, 1, pad=(p, p), out=out))
    dy = torch.randn(n, h, w, cv.round8(cout), device=dev).to(torch.bfloat16)
    dx = torch.empty(n, h, w, cin, device=dev, dtype=torch.bfloat16)
    t_d = timeit(lambda: cv.dgrad(dy, wd, cin, cout, k, k, (h, w), 1, pad=(p, p), out=dx))
    dw = torch.zeros(k * k, cout, cin, device=dev)
    t_w = timeit(lambda: cv.wgrad(x, dy, dw, cin, cout, k, k, 1, pad=(p, p)))
    xc = x.permute(0, 3, 1, 2)
    wc = wt.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    t_c = timeit(lambda: F.conv2d(xc, wc, padding=p))
    dyc = dy[..., :cout].permute(0, 3, 1, 2) if cv.round8(cout) == cout else dy[..., :cout].contiguous().permute(0, 3, 1, 2)
    t_cd = timeit(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [1, 1], [p, p], [1, 1], False, [0, 0], 1, [True, False, False]))
    t_cw = timeit(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [1, 1], [p, p], [1, 1], False, [0, 0], 1, [False, True, False]))
    row = dict(name=name, gflop=flops / 1e9, fprop_ms=t_f, fprop_tflops=flops / t_f / 1e9, dgrad_ms=t_d,
               dgrad_tflops=flops / t_d / 1e9, wgrad_ms=t_w, wgrad_tflops=flops / t_w / 1e9, cudnn_fprop_ms=t_c,
               cudnn_fprop_tflops=flops / t_c / 1e9, cudnn_dgrad_ms=t_cd, cudnn_wgrad_ms=t_cw)
    rows.append(row)
    print(json.dumps(row), flush=True)
