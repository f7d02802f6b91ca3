"""Run one wgrad shape a few times (for ncu captures).  Usage: one_wgrad.py n h w cin cout k"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.ops import conv as cv
n, h, w, cin, cout, k = [int(v) for v in sys.argv[1:7]]
x = torch.randn(n, h, w, cin, device="cuda").to(torch.bfloat16)
dy = torch.randn(n, h, w, cv.round8(cout), device="cuda").to(torch.bfloat16)
dw = torch.zeros(k * k, cout, cin, device="cuda")
for _ in range(4):
    cv.wgrad(x, dy, dw, cin, cout, k, k, 1, pad=(k // 2, k // 2))
torch.cuda.synchronize()
print("done")
