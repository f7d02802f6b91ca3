import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.ops import conv as cv
n, h, w, cin, cout, k = [int(v) for v in sys.argv[1:7]]
res = len(sys.argv) > 7 and sys.argv[7] == "1"
x = torch.randn(n, h, w, cin, device="cuda").to(torch.bfloat16)
wt = torch.randn(cout, cin, k, k, device="cuda") * 0.02
wf, wd = cv.pack_working(cv.pack_master(wt))
out = torch.empty(n, h, w, cv.round8(cout), device="cuda", dtype=torch.bfloat16)
add = torch.randn_like(out) if res else None
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for i in range(8):
    flush.zero_()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); cv.fprop(x, wf, cin, cout, k, k, 1, pad=(k // 2, k // 2), out=out, relu=True, add=add); e.record()
    torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
print(os.environ.get("B200_CONV_DEBUG", "0"), sys.argv[1:], "median ms", sorted(ts)[4])
