"""VGG backbone on the native kernels (engine/vgg_native.py) vs the PyTorch spec model + autograd in fp32."""
import copy

import pytest
import torch
import torch.nn.functional as F

from batchai_retinanet_horovod_coco_b200 import models
from test_gpu_engine import _cos, _inputs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(2, 8, 12, 16), (3, 9, 13, 64), (1, 2, 3, 8)])
def test_maxpool2_matches_torch(shape):
    from batchai_retinanet_horovod_coco_b200.ops import native
    n, h, w, c = shape
    torch.manual_seed(0)
    x = torch.relu(torch.randn(n, h, w, c, device="cuda")).to(torch.bfloat16)
    x[0, :2, :2] = 0                                   # an all-zero window passes no gradient (ReLU'(0) = 0)
    y = torch.empty(n, h // 2, w // 2, c, device="cuda", dtype=torch.bfloat16)
    native.maxpool2_fwd(x, y)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.max_pool2d(xr, 2, 2)
    assert torch.equal(y.float(), yr.permute(0, 2, 3, 1))
    dy = torch.randn_like(y)
    dx = torch.full_like(x, 7.0)
    native.maxpool2_relu_bwd(x, dy, dx)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    want = xr.grad.permute(0, 2, 3, 1) * (x.float() > 0)
    # ties: both pick the first maximum in scan order; compare where the maximum is unique, and totals everywhere
    assert torch.allclose(dx.float().sum(dim=(1, 2)), want.sum(dim=(1, 2)), atol=1e-2)
    win = F.unfold(x.float().permute(0, 3, 1, 2)[:, :, :h // 2 * 2, :w // 2 * 2].reshape(n * c, 1, h // 2 * 2, w // 2 * 2), 2, stride=2)
    assert (dx.float()[:, h // 2 * 2:] == 0).all() and (dx.float()[:, :, w // 2 * 2:] == 0).all()
    uniq = (win == win.max(dim=1, keepdim=True).values).sum(dim=1) == 1
    if uniq.all():
        assert torch.equal(dx.float(), want)


@pytest.mark.parametrize("name,hw", [("vgg16", (160, 224)), ("vgg19", (166, 230))])
def test_vgg_native_matches_autograd(name, hw):
    from batchai_retinanet_horovod_coco_b200.engine import create_trainer
    from batchai_retinanet_horovod_coco_b200.engine.library import LibraryTrainer
    from batchai_retinanet_horovod_coco_b200.engine.vgg_native import VGGTrainer
    torch.manual_seed(0)
    B, (H, W) = 2, hw
    model = models.backbone(name).retinanet(80)
    with torch.no_grad():
        for sm in (model.regression_submodel, model.classification_submodel):
            for c in list(sm.tower) + [sm.out]:
                c.weight.normal_(0, 0.03)
        for c in model.backbone.keras_layers().values():
            c.bias.normal_(0, 0.05)
    ref_model = copy.deepcopy(model)
    img, gt, cnt, sz = _inputs(B, H, W)
    assert isinstance(create_trainer(copy.deepcopy(model), "cuda", batch=B, image_hw=(H, W)), VGGTrainer)
    vt = VGGTrainer(model=model, device="cuda", batch=B, image_hw=(H, W), shape_bucket=1)
    loss_v, _ = vt.forward_loss(img, gt, cnt, sz)
    torch.cuda.synchronize()
    grads = vt.gradients_by_layer()

    lib = LibraryTrainer(model=ref_model, device="cuda", dtype=torch.float32)
    total, lc, lr_ = lib.loss(img.cuda(), gt.cuda(), cnt.cuda(), sz.cuda())
    total.backward()
    ref = torch.stack([total, lr_, lc]).detach()
    rel = ((loss_v - ref).abs() / ref.abs().clamp(min=1e-6)).max().item()
    assert rel < 5e-2, f"loss mismatch native {loss_v.tolist()} vs ref {ref.tolist()}"
    bad = []
    for lname, mod in ref_model.keras_layers().items():
        gw, gb = grads[lname]
        cw = _cos(gw, mod.weight.grad)
        ratio = float(gw.norm() / (mod.weight.grad.norm() + 1e-30))
        if cw < 0.97 or not (0.9 < ratio < 1.1):
            bad.append((lname, "w", round(cw, 4), round(ratio, 3)))
        if _cos(gb, mod.bias.grad) < 0.97:
            bad.append((lname, "b", round(_cos(gb, mod.bias.grad), 4)))
    assert not bad, f"gradient mismatches: {bad[:20]} ({len(bad)} total)"
