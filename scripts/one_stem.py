"""Run the stem forward and weight-gradient kernels at full size (for ncu captures).  Usage: one_stem.py [n h w]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.ops import native
n, h, w = [int(v) for v in sys.argv[1:4]] if len(sys.argv) >= 4 else (8, 800, 1333)
x4 = torch.randn(n, h, w, 4, device="cuda").to(torch.bfloat16)
wp = native.pack_stem_weight(torch.randn(49, 64, 3, device="cuda") * 0.05)
shift = torch.zeros(64, device="cuda")
scale = torch.ones(64, device="cuda")
y = torch.empty(n, (h + 1) // 2, (w + 1) // 2, 64, device="cuda", dtype=torch.bfloat16)
dw = torch.zeros(49, 64, 3, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
for it in range(4):
    ev[0].record()
    native.stem_conv(x4, wp, shift, y)
    ev[1].record()
    native.stem_wgrad(x4, y, scale, dw)
    ev[2].record()
torch.cuda.synchronize()
print("stem_conv ms", ev[0].elapsed_time(ev[1]), "stem_wgrad ms", ev[1].elapsed_time(ev[2]))
