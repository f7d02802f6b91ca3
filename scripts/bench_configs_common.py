"""Shared helpers of the one-GPU bench scripts: synthetic batch and CUDA-event timing."""
import torch

dev = torch.device("cuda", 0)


def batch(B, H, W, G=12, seed=0):
    g = torch.Generator().manual_seed(seed)
    img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g).to(dev)
    xy = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.7, H * 0.7])
    wh = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.3, H * 0.3]) + 24
    gt = torch.cat([xy, xy + wh, torch.randint(0, 80, (B, G, 1), generator=g).float()], dim=2).to(dev)
    return img, gt, torch.full((B,), G, dtype=torch.int32, device=dev), torch.tensor([[H, W]] * B, dtype=torch.int32, device=dev)


def timeit(fn, steps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / steps
