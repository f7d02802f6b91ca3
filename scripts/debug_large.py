import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from batchai_retinanet_horovod_coco_b200.ops import conv as cv

def ref_conv(x, c, wt, pad):
    return F.conv2d(x[..., :c].float().permute(0, 3, 1, 2), wt.to(torch.bfloat16).float(), padding=pad).permute(0, 2, 3, 1)

torch.manual_seed(0)
for (n, h, w, cin, cout, k) in [(8, 100, 167, 256, 256, 3), (8, 200, 334, 64, 64, 3), (8, 200, 334, 64, 256, 1), (2, 50, 84, 1024, 256, 1), (8, 100, 167, 256, 720, 3)]:
    x = torch.randn(n, h, w, cin, device="cuda").to(torch.bfloat16)
    wt = torch.randn(cout, cin, k, k, device="cuda") / (cin * k * k) ** 0.5
    wf, wd = cv.pack_working(cv.pack_master(wt))
    y = cv.fprop(x, wf, cin, cout, k, k, 1, pad=(k // 2, k // 2))
    torch.cuda.synchronize()
    ref = ref_conv(x, cin, wt, k // 2)
    err = (y[..., :cout].float() - ref).abs()
    print("fprop", (n, h, w, cin, cout, k), "max err", err.max().item(), "ref max", ref.abs().max().item(),
          "bad frac", (err > 0.05).float().mean().item())
    if err.max() > 0.1:
        bad = (err > 0.1).nonzero()
        print("  first bad idx", bad[:5].tolist(), " last bad", bad[-3:].tolist())
        # which images / rows are bad
        print("  bad per image", [(err[i] > 0.1).float().mean().item() for i in range(n)])
    dy = torch.randn(n, h, w, cv.round8(cout), device="cuda").to(torch.bfloat16)
    dw = torch.zeros(k * k, cout, cin, device="cuda")
    cv.wgrad(x, dy, dw, cin, cout, k, k, 1, pad=(k // 2, k // 2))
    wz = torch.zeros(cout, cin, k, k, device="cuda", requires_grad=True)
    yy = F.conv2d(x.float().permute(0, 3, 1, 2), wz, padding=k // 2)
    yy.backward(dy[..., :cout].float().permute(0, 3, 1, 2))
    refw = cv.pack_master(wz.grad)
    errw = (dw - refw).abs().max().item()
    print("wgrad", "max err", errw, "ref max", refw.abs().max().item())
