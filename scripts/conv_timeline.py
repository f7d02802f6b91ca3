"""In-kernel timeline of one conv launch (b200_conv_set_trace): where a CTA's first tile spends its cycles.
Needs the traced build: B200_CONV_TRACE=1 python build_native.py (the default build compiles the stamps out).
Usage: python scripts/conv_timeline.py [n h w cin cout k]   (default: res4 1x1 1024->256 at 8x50x84)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.ops import conv as cv, native

a = [int(v) for v in sys.argv[1:7]] or [8, 50, 84, 1024, 256, 1]
n, h, w, cin, cout, k = a
x = torch.randn(n, h, w, cin, device="cuda").to(torch.bfloat16)
wt = torch.randn(cout, cin, k, k, device="cuda") * 0.02
wf, _ = cv.pack_working(cv.pack_master(wt))
out = torch.empty(n, h, w, cv.round8(cout), device="cuda", dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
buf = torch.zeros(160 * 16, dtype=torch.int64, device="cuda")
run = lambda: cv.fprop(x, wf, cin, cout, k, k, 1, pad=(k // 2, k // 2), out=out)
for cold in (False, True):
    for _ in range(3):
        run()
    if cold:
        flush.zero_()
    torch.cuda.synchronize()
    native.conv_set_trace(buf)
    buf.zero_()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); run(); e.record()
    torch.cuda.synchronize()
    native.conv_set_trace(None)
    t = buf.view(-1, 16).cpu()
    t = t[t[:, 0] > 0]
    g0 = t[:, 0].min().item()
    names = {2: "set-up done", 3: "first A box issued", 4: "first A box landed", 5: "first B tile landed", 13: "tile0 last A box landed",
             6: "tile0 last MMA issued", 7: "tile0 accumulator complete", 8: "tile0 epilogue done", 9: "last tile last MMA issued",
             14: "last tile accumulator complete", 10: "last store read out", 11: "exit"}
    print(f"\n{'cold L2' if cold else 'warm'}: {n}x{h}x{w} {cin}->{cout} {k}x{k}: event time {s.elapsed_time(e) * 1e3:.1f} us, {t.shape[0]} CTAs, "
          f"CTA start spread {(t[:, 0].max().item() - g0) / 1e3:.2f} us, kernel span by globaltimer {(t[:, 12].max().item() - g0) / 1e3:.2f} us")
    print("cycles since CTA entry (median / min / max over CTAs):")
    for i in (2, 3, 4, 5, 13, 6, 7, 8, 9, 14, 10, 11):
        d = (t[:, i] - t[:, 1])[t[:, i] > 0].float()
        if d.numel():
            print(f"  {names[i]:32s} {d.median().item():9.0f} {d.min().item():9.0f} {d.max().item():9.0f}   ({d.numel()} CTAs)")
