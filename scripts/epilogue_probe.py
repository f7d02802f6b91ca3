"""How much do the epilogue operands cost?  1x1 conv 64 -> 256 at 8x200x334 (res2 branch2c geometry) with no
operand, + residual, + mask, + both; and the mirrored 256 -> 64 reduction.  Device-timed, 20 launches each."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.ops import native
bf = torch.bfloat16
N, H, W = 8, 200, 334
x64 = torch.randn(N, H, W, 64, device="cuda").to(bf)
x256 = torch.randn(N, H, W, 256, device="cuda").to(bf)
w_up = (torch.randn(1, 256, 64, device="cuda") * 0.1).to(bf)
w_dn = (torch.randn(1, 64, 256, device="cuda") * 0.05).to(bf)
res = torch.randn(N, H, W, 256, device="cuda").to(bf)
msk = torch.randn(N, H, W, 256, device="cuda").to(bf)
y256 = torch.empty_like(res)
y64 = torch.empty_like(x64)
shift = torch.zeros(256, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def t(fn, label, bytes_moved):
    for _ in range(3):
        fn()
    ms = []
    for _ in range(10):
        flush.zero_()                       # evict L2 between launches
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ms.append(s.elapsed_time(e))
    m = sorted(ms)[len(ms) // 2]
    print(f"{label:34s} {m * 1e3:7.1f} us  {bytes_moved / m / 1e9:6.2f} TB/s")


px = N * H * W
up = lambda add, mask, sh, relu: native.conv_raw(0, [(x64, y256, add, mask)], w_up, 64, 64, 256, 256, 1, 1, 1, 0, 0, shift=sh, relu=relu)
t(lambda: up(None, None, None, False), "64->256 plain", px * (64 + 256) * 2)
t(lambda: up(None, None, shift, True), "64->256 + shift + relu", px * (64 + 256) * 2)
t(lambda: up(res, None, shift, True), "64->256 + shift + residual + relu", px * (64 + 512) * 2)
t(lambda: up(None, msk, None, False), "64->256 + mask", px * (64 + 512) * 2)
t(lambda: up(res, msk, None, False), "64->256 + residual + mask", px * (64 + 768) * 2)
dn = lambda mask: native.conv_raw(0, [(x256, y64, None, mask)], w_dn, 256, 256, 64, 64, 1, 1, 1, 0, 0)
t(lambda: dn(None), "256->64 plain", px * (256 + 64) * 2)
t(lambda: dn(x64), "256->64 + mask", px * (256 + 128) * 2)
t(lambda: y256.copy_(res), "torch copy 256ch (reference)", px * 512 * 2)
