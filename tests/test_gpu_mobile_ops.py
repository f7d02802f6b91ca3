"""Depthwise conv, batch-statistics BN (+ReLU/ReLU6), 2x2 average pool and the small-Cin filter gradient vs plain
PyTorch fp32 references of the same ops; plus the tensor-core wgrad with Cin that is not a multiple of 64."""
import pytest
import torch
import torch.nn.functional as F

from batchai_retinanet_horovod_coco_b200.models.resnet import tf_same_pad
from batchai_retinanet_horovod_coco_b200.ops import conv as cv
from batchai_retinanet_horovod_coco_b200.ops import native

pytestmark = pytest.mark.gpu
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
DEV = "cuda"
BF = torch.bfloat16


def _close(a, b, rel, ab_rms):
    a, b = a.float(), b.float()
    rms = b.pow(2).mean().sqrt().item() + 1e-12
    bad = (a - b).abs() > rel * b.abs() + ab_rms * rms
    assert not bool(bad.any()), f"{int(bad.sum())}/{bad.numel()} out of bound, worst {(a - b).abs().max().item():.4g} rms {rms:.4g}"


@pytest.mark.parametrize("n,h,w,c,stride", [(2, 19, 23, 32, 1), (2, 20, 31, 64, 2), (1, 7, 9, 512, 1), (3, 16, 16, 24, 2)])
def test_depthwise(n, h, w, c, stride):
    torch.manual_seed(0)
    x = torch.randn(n, h, w, c, device=DEV).to(BF)
    wt = torch.randn(c, 1, 3, 3, device=DEV) * 0.3
    ph, pw = tf_same_pad(h, 3, stride), tf_same_pad(w, 3, stride)
    oh, ow = -(-h // stride), -(-w // stride)
    xr = F.pad(x.float().permute(0, 3, 1, 2), (pw[0], pw[1], ph[0], ph[1])).requires_grad_(True)
    wr = wt.clone().requires_grad_(True)
    yr = F.conv2d(xr, wr, stride=stride, groups=c)
    y = torch.empty(n, oh, ow, c, device=DEV, dtype=BF)
    native.dwconv3x3(0, x, wt.reshape(c, 9), y, (h, w), (oh, ow), stride, (ph[0], pw[0]))
    _close(y, yr.permute(0, 2, 3, 1), 2 ** -7, 2e-3)
    dy = torch.randn(n, oh, ow, c, device=DEV).to(BF)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dx = torch.empty_like(x)
    native.dwconv3x3(1, dy, wt.reshape(c, 9), dx, (h, w), (oh, ow), stride, (ph[0], pw[0]))
    dxr = xr.grad[:, :, ph[0]:ph[0] + h, pw[0]:pw[0] + w].permute(0, 2, 3, 1)
    _close(dx, dxr, 2 ** -7, 2e-3)
    dw = torch.zeros(c, 9, device=DEV)
    native.dwconv3x3(2, x, dy, dw, (h, w), (oh, ow), stride, (ph[0], pw[0]))
    _close(dw, wr.grad.reshape(c, 9), 1e-3, 1e-3)


@pytest.mark.parametrize("act", [0, 1, 2])
def test_batchnorm_batch_statistics(act):
    torch.manual_seed(1)
    n, h, w, c = 3, 17, 21, 72
    x = (torch.randn(n, h, w, c, device=DEV) * 2 + 0.5).to(BF)
    gamma = torch.rand(c, device=DEV) + 0.5
    beta = torch.randn(c, device=DEV) * 2 + (2.0 if act == 2 else 0.0)
    eps = 1e-3
    sums = torch.zeros(2 * c, device=DEV)
    mean, var = torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    mm, mv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    native.bn_stats(x, c, sums, mean, var, mm, mv, 0.99)
    xf = x.float().reshape(-1, c)
    _close(mean, xf.mean(0), 1e-4, 1e-4)
    _close(var, xf.var(0, unbiased=False), 1e-3, 1e-4)
    _close(mm, 0.01 * xf.mean(0), 1e-3, 1e-4)
    _close(mv, 0.99 + 0.01 * xf.var(0, unbiased=False), 1e-3, 1e-4)
    assert float(sums.abs().max()) == 0.0
    y = torch.empty_like(x)
    native.bn_act_fwd(x, mean, var, gamma, beta, eps, y, c, act)
    xr = x.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    z = F.batch_norm(xr.reshape(-1, c), None, None, gr, br, True, 0.0, eps).reshape(n, h, w, c)
    yr = z if act == 0 else (F.relu(z) if act == 1 else F.relu6(z))
    _close(y, yr, 2 ** -7, 3e-3)
    dy = torch.randn(n, h, w, c, device=DEV).to(BF)
    yr.backward(dy.float())
    dx = torch.empty_like(x)
    dg, db = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    native.bn_act_bwd(x, dy, mean, var, gamma, beta, eps, sums, c, act, dx, dg, db)
    _close(db, br.grad, 2e-3, 2e-3)
    _close(dg, gr.grad, 2e-3, 2e-3)
    _close(dx, xr.grad, 2 ** -6, 5e-3)
    # accumulate into a slice of a wider buffer
    wide = torch.ones(n, h, w, c + 24, device=DEV, dtype=BF)
    native.bn_act_bwd(x, dy, mean, var, gamma, beta, eps, sums, c, act, wide, None, None, accumulate=True)
    _close(wide[..., :c], xr.grad + 1.0, 2 ** -6, 5e-3)
    assert bool((wide[..., c:] == 1).all())


def test_avgpool2():
    x = torch.randn(2, 9, 13, 40, device=DEV).to(BF)
    y = torch.empty(2, 4, 6, 40, device=DEV, dtype=BF)
    native.avgpool2(x, y)
    _close(y, F.avg_pool2d(x.float().permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1), 2 ** -7, 2e-3)
    dy = torch.randn_like(y.float()).to(BF)
    dx = torch.empty_like(x)
    native.avgpool2(dy, dx, backward=True)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    F.avg_pool2d(xr, 2).backward(dy.float().permute(0, 3, 1, 2))
    _close(dx, xr.grad.permute(0, 2, 3, 1), 2 ** -7, 1e-3)


@pytest.mark.parametrize("cout,stride", [(32, 2), (64, 1)])
def test_wgrad_small_first_layer(cout, stride):
    n, h, w = 2, 21, 30
    x = torch.zeros(n, h, w, 8, device=DEV, dtype=BF)
    x[..., :3] = torch.randn(n, h, w, 3, device=DEV).to(BF)
    ph, pw = tf_same_pad(h, 3, stride), tf_same_pad(w, 3, stride)
    oh, ow = -(-h // stride), -(-w // stride)
    dy = torch.randn(n, oh, ow, cout, device=DEV).to(BF)
    dw = torch.zeros(9, cout, 3, device=DEV)
    native.wgrad_small(x, dy, dw, 3, cout, 3, 3, stride, (ph[0], pw[0]))
    wz = torch.zeros(cout, 3, 3, 3, device=DEV, requires_grad=True)
    xr = F.pad(x[..., :3].float().permute(0, 3, 1, 2), (pw[0], pw[1], ph[0], ph[1]))
    F.conv2d(xr, wz, stride=stride).backward(dy.float().permute(0, 3, 1, 2))
    _close(dw, cv.pack_master(wz.grad), 1e-3, 1e-3)


@pytest.mark.parametrize("cin,cout,k", [(32, 64, 1), (96, 128, 1), (160, 32, 3), (24, 48, 1)])
def test_tensor_core_wgrad_partial_channel_chunk(cin, cout, k):
    """Cin that does not fill its last 64-channel chunk (MobileNet 32 / 24.., DenseNet 64 + 32 j): TMA zero-fills it."""
    n, h, w = 2, 18, 25
    x = torch.randn(n, h, w, cin, device=DEV).to(BF)
    dy = torch.randn(n, h, w, cout, device=DEV).to(BF)
    dw = torch.zeros(k * k, cout, cin, device=DEV)
    cv.wgrad(x, dy, dw, cin, cout, k, k, 1, pad=(k // 2, k // 2))
    wz = torch.zeros(cout, cin, k, k, device=DEV, requires_grad=True)
    F.conv2d(x.float().permute(0, 3, 1, 2), wz, padding=k // 2).backward(dy.float().permute(0, 3, 1, 2))
    _close(dw, cv.pack_master(wz.grad), 1e-3, 1e-3)
