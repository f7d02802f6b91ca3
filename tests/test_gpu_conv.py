"""tcgen05 convolution kernels vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

from batchai_retinanet_horovod_coco_b200.ops import conv as cv
from batchai_retinanet_horovod_coco_b200.ops import native

pytestmark = pytest.mark.gpu

# the fp32 oracles must be fp32: PyTorch would otherwise run its convolutions in TF32 (10-bit mantissa)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def _mk(n, h, w, c, ld=None, scale=1.0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    ld = ld or cv.round8(c)
    x = torch.zeros(n, h, w, ld, device="cuda", dtype=torch.bfloat16)
    x[..., :c] = (torch.randn(n, h, w, c, device="cuda", generator=g) * scale).to(torch.bfloat16)
    return x


def _ref_conv(x_nhwc, c, w_oihw, stride, pads):
    """pads = (top, bottom, left, right)"""
    x = x_nhwc[..., :c].float().permute(0, 3, 1, 2)
    x = F.pad(x, (pads[2], pads[3], pads[0], pads[1]))
    return F.conv2d(x, w_oihw.to(torch.bfloat16).float(), stride=stride)


def _close(a, b, tol=None):
    """Per-element bound.  bf16 outputs: one rounding of an fp32 result is at most 2^-8 relative, so 2^-7 |ref| plus a
    small absolute term tied to the tensor's RMS (accumulation order, results that cancel to ~0).  fp32 outputs
    (weight gradients): 1e-3 |ref| + 5e-4 rms.  A dropped K slice, a wrong channel tail or one mis-swizzled 16-byte
    piece moves single elements by ~rms and fails; the old max-abs-vs-max-abs test could not see those."""
    a32, b32 = a.float(), b.float()
    rms = b32.pow(2).mean().sqrt().item() + 1e-12
    if a.dtype == torch.bfloat16:
        rel, ab = 2.0 ** -7, 2e-3 * rms
    else:
        rel, ab = 1e-3, 5e-4 * rms
    err = (a32 - b32).abs()
    bound = rel * b32.abs() + ab
    bad = err > bound
    assert not bool(bad.any()), (f"{int(bad.sum())} of {bad.numel()} elements out of bound; worst err "
                                 f"{err.max().item():.4g} (rms {rms:.4g}) at {tuple(torch.nonzero(bad)[0].tolist())}")
    return (err / (b32.abs() + rms)).max().item()


def test_fp64_spot_check_of_the_oracle():
    """The fp32 PyTorch oracle itself against float64 on the CPU for one shape family (TF32 is disabled above)."""
    n, h, w, cin, cout, k = 1, 9, 11, 256, 64, 3
    x = _mk(n, h, w, cin)
    wt = torch.randn(cout, cin, k, k, device="cuda") * 0.02
    ref32 = _ref_conv(x, cin, wt, 1, (1, 1, 1, 1))
    ref64 = F.conv2d(F.pad(x[..., :cin].double().cpu().permute(0, 3, 1, 2), (1, 1, 1, 1)),
                     wt.to(torch.bfloat16).double().cpu())
    assert (ref32.double().cpu() - ref64).abs().max().item() <= 1e-4 * ref64.abs().max().item()
    wf, _ = cv.pack_working(cv.pack_master(wt))
    y = cv.fprop(x, wf, cin, cout, k, k, 1, pad=(1, 1))
    _close(y[..., :cout], ref64.permute(0, 2, 3, 1).float().cuda())


CASES = [
    # n, h, w, cin, cout, k, stride
    (2, 16, 16, 64, 64, 1, 1),
    (2, 13, 21, 256, 512, 1, 1),
    (1, 25, 42, 256, 256, 3, 1),
    (2, 7, 11, 256, 720, 3, 1),
    (2, 9, 10, 256, 36, 3, 1),
    (2, 26, 42, 256, 128, 1, 2),
    (1, 25, 42, 128, 256, 3, 2),
    (1, 32, 32, 2048, 256, 3, 2),
    # streamed weights -> CTA pairs sharing the weight ring by TMA multicast: odd tile counts (the odd CTA of the last
    # pair only forwards weight tiles), more tiles than CTAs, several N tiles
    (1, 24, 40, 256, 256, 3, 1),
    (8, 50, 84, 1024, 256, 1, 1),
    (5, 50, 75, 256, 1024, 1, 1),
]


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride", CASES)
def test_fprop(n, h, w, cin, cout, k, stride):
    torch.manual_seed(1)
    x = _mk(n, h, w, cin)
    wt = torch.randn(cout, cin, k, k, device="cuda") * (1.0 / (cin * k * k) ** 0.5)
    if stride == 1:
        pb = (k // 2, k // 2)
        pads = (k // 2, k // 2, k // 2, k // 2)
        ptot = (2 * pb[0], 2 * pb[1])
    else:
        from batchai_retinanet_horovod_coco_b200.models.resnet import tf_same_pad
        ph, pw = tf_same_pad(h, k, 2), tf_same_pad(w, k, 2)
        pb, pads, ptot = (ph[0], pw[0]), (ph[0], ph[1], pw[0], pw[1]), (sum(ph), sum(pw))
    wf, _ = cv.pack_working(cv.pack_master(wt))
    y = cv.fprop(x, wf, cin, cout, k, k, stride, pad=pb, pad_total=ptot)
    torch.cuda.synchronize()
    ref = _ref_conv(x, cin, wt, stride, pads).permute(0, 2, 3, 1)
    assert y.shape[:3] == ref.shape[:3]
    _close(y[..., :cout], ref)


@pytest.mark.parametrize("case", [(1, 25, 42, 256, 256, 3, 1), (2, 7, 11, 256, 720, 3, 1), (2, 13, 21, 256, 512, 1, 1),
                                  (1, 24, 40, 256, 256, 3, 1)])
def test_fprop_and_dgrad_with_wide_tiles_on_small_problems(case, monkeypatch):
    """Small problems split their output channels into 128- / 64-column tiles by default (more CTAs); the 256-column
    path they would otherwise take (resident or streamed weights, four chunks per tile) stays covered here."""
    monkeypatch.setenv("B200_CONV_SPLIT_N", "0")
    test_fprop(*case)
    test_dgrad(*case)


def test_fprop_epilogue():
    n, h, w, cin, cout = 2, 20, 30, 128, 256
    x = _mk(n, h, w, cin)
    wt = torch.randn(cout, cin, 3, 3, device="cuda") * 0.03
    shift = torch.randn(cout, device="cuda")
    res = _mk(n, h, w, cout, seed=3)
    wf, _ = cv.pack_working(cv.pack_master(wt))
    y = cv.fprop(x, wf, cin, cout, 3, 3, 1, pad=(1, 1), shift=shift, relu=True, add=res)
    ref = _ref_conv(x, cin, wt, 1, (1, 1, 1, 1)).permute(0, 2, 3, 1) + shift + res.float()
    _close(y, ref.clamp(min=0))
    y2 = cv.fprop(x, wf, cin, cout, 3, 3, 1, pad=(1, 1), shift=shift)
    _close(y2, _ref_conv(x, cin, wt, 1, (1, 1, 1, 1)).permute(0, 2, 3, 1) + shift)


@pytest.mark.parametrize("cin,cout,k,add,mask,post", [
    (64, 256, 1, True, False, False),      # res2 'c' conv: one K stage, four 64-column chunks, residual in place
    (64, 64, 3, False, True, False),       # one chunk per tile: the two epilogue groups alternate tiles
    (256, 1024, 1, True, True, False),     # four N tiles, residual + mask (two slot positions per chunk)
    (128, 512, 1, True, True, True),       # add applied after the mask
    (512, 192, 1, False, False, False),    # three chunks per tile: chunk -> group mapping flips every tile
    (256, 720, 3, False, True, False),     # N tiles of 256 with a 208-column tail
    (128, 40, 1, True, True, False),       # channel count below one chunk
])
def test_epilogue_variants(cin, cout, k, add, mask, post):
    """Every operand combination of the TMA-fed epilogue at shapes with partial tiles in both pixel dimensions."""
    n, h, w = 3, 37, 53
    x = _mk(n, h, w, cin)
    wt = torch.randn(cout, cin, k, k, device="cuda") * (1.0 / (cin * k * k) ** 0.5)
    shift = torch.randn(cout, device="cuda")
    wf, _ = cv.pack_working(cv.pack_master(wt))
    ldo = cv.round8(cout)
    res = _mk(n, h, w, cout, seed=11) if add else None
    msk = _mk(n, h, w, cout, seed=12) if mask else None
    out = torch.full((n, h, w, ldo), 7.0, device="cuda", dtype=torch.bfloat16)
    native.conv_raw(0, [(x, out, res, msk)], wf, cin, x.shape[-1], cout, ldo, k, k, 1, k // 2, k // 2, shift=shift,
                    relu=True, add_after_mask=post)
    ref = _ref_conv(x, cin, wt, 1, (k // 2,) * 4).permute(0, 2, 3, 1) + shift
    if add and not post:
        ref = ref + res[..., :cout].float()
    if mask:
        ref = ref * (msk[..., :cout].float() > 0)
    if add and post:
        ref = ref + res[..., :cout].float()
    _close(out[..., :cout], ref.clamp(min=0))
    if ldo != cout:
        assert bool((out[..., cout:] == 0).all()), "channel padding must stay zero"


def test_epilogue_in_place_residual():
    """out aliases add (the engine's  g <- mask(dgrad + g)  update)."""
    n, h, w, cin, cout = 2, 20, 30, 64, 256
    wt = torch.randn(cout, cin, 1, 1, device="cuda") * 0.1
    dy = _mk(n, h, w, cout)
    _, wd = cv.pack_working(cv.pack_master(wt))
    g = _mk(n, h, w, cin, seed=4)
    mask = _mk(n, h, w, cin, seed=5)
    ref = (_ref_dgrad(dy, cout, wt, 1, (0, 0, 0, 0), (h, w)) + g.float()) * (mask.float() > 0)
    native.conv_raw(1, [(dy, g, g, mask)], wd, cout, dy.shape[-1], cin, g.shape[-1], 1, 1, 1, 0, 0)
    _close(g, ref)


def test_dgrad_column_sums():
    """The store warp's column sums (bias gradient of the layer whose dY this output is) equal the sum of the stored
    bf16 tensor over pixels: grouped launch into one shared vector, stride-2 parity launches, several N tiles."""
    shapes = [(2, 25, 42), (2, 13, 21), (2, 7, 11)]
    wt = torch.randn(256, 256, 3, 3, device="cuda") * 0.02
    _, wd = cv.pack_working(cv.pack_master(wt))
    dys = [_mk(n, h, w, 256, seed=i) for i, (n, h, w) in enumerate(shapes)]
    masks = [_mk(n, h, w, 256, seed=20 + i) for i, (n, h, w) in enumerate(shapes)]
    outs = [torch.empty_like(d) for d in dys]
    db = torch.full((256,), 3.0, device="cuda")
    native.conv_raw(1, [(d, o, None, m, db) for d, o, m in zip(dys, outs, masks)], wd, 256, 256, 256, 256, 3, 3, 1, 1, 1)
    want = 3.0 + sum(o.float().sum(dim=(0, 1, 2)) for o in outs)
    assert torch.allclose(db, want, rtol=1e-4, atol=1e-3 * float(want.abs().max())), (db - want).abs().max()
    # 1x1, 64 -> 1024 channels (four N tiles), residual + mask, partial tiles
    n, h, w, cin, cout = 2, 19, 23, 1024, 64
    wt = torch.randn(cout, cin, 1, 1, device="cuda") * 0.1
    _, wd = cv.pack_working(cv.pack_master(wt))
    dy = _mk(n, h, w, cout)
    add, mask = _mk(n, h, w, cin, seed=3), _mk(n, h, w, cin, seed=4)
    out = torch.empty(n, h, w, cin, device="cuda", dtype=torch.bfloat16)
    db = torch.zeros(cin, device="cuda")
    native.conv_raw(1, [(dy, out, add, mask, db)], wd, cout, dy.shape[-1], cin, cin, 1, 1, 1, 0, 0)
    want = out.float().sum(dim=(0, 1, 2))
    assert torch.allclose(db, want, rtol=1e-4, atol=1e-3 * float(want.abs().max()))
    # 3x3 stride 2 (four parity launches into one vector)
    from batchai_retinanet_horovod_coco_b200.models.resnet import tf_same_pad
    h, w, cin, cout = 25, 42, 256, 256
    ph, pw = tf_same_pad(h, 3, 2), tf_same_pad(w, 3, 2)
    wt = torch.randn(cout, cin, 3, 3, device="cuda") * 0.02
    _, wd = cv.pack_working(cv.pack_master(wt))
    dy = _mk(1, 13, 21, cout)
    out = torch.zeros(1, h, w, cin, device="cuda", dtype=torch.bfloat16)
    db = torch.zeros(cin, device="cuda")
    for py in range(2):
        for px in range(2):
            native.conv_raw(1, [(dy, out, None, None, db)], wd, cout, cout, cin, cin, 3, 3, 2, ph[0], pw[0], par=(py, px))
    want = out.float().sum(dim=(0, 1, 2))
    assert torch.allclose(db, want, rtol=1e-4, atol=1e-3 * float(want.abs().max()))


def test_dgrad_group_masks():
    shapes = [(2, 25, 42), (2, 13, 21), (2, 7, 11), (2, 4, 6), (2, 2, 3)]
    wt = torch.randn(256, 256, 3, 3, device="cuda") * 0.02
    _, wd = cv.pack_working(cv.pack_master(wt))
    dys = [_mk(n, h, w, 256, seed=i) for i, (n, h, w) in enumerate(shapes)]
    masks = [_mk(n, h, w, 256, seed=20 + i) for i, (n, h, w) in enumerate(shapes)]
    outs = [torch.empty_like(d) for d in dys]
    native.conv_raw(1, [(d, o, None, m) for d, o, m in zip(dys, outs, masks)], wd, 256, 256, 256, 256, 3, 3, 1, 1, 1)
    for (n, h, w), d, o, m in zip(shapes, dys, outs, masks):
        ref = _ref_dgrad(d, 256, wt, 1, (1, 1, 1, 1), (h, w)) * (m.float() > 0)
        _close(o, ref)


def test_fprop_group():
    shapes = [(2, 25, 42), (2, 13, 21), (2, 7, 11), (2, 4, 6), (2, 2, 3)]
    xs = [_mk(n, h, w, 256, seed=i) for i, (n, h, w) in enumerate(shapes)]
    wt = torch.randn(256, 256, 3, 3, device="cuda") * 0.02
    b = torch.randn(256, device="cuda")
    wf, _ = cv.pack_working(cv.pack_master(wt))
    ys = cv.fprop_group(xs, wf, 256, 256, 3, 3, shift=b, relu=True)
    for x, y in zip(xs, ys):
        ref = (_ref_conv(x, 256, wt, 1, (1, 1, 1, 1)).permute(0, 2, 3, 1) + b).clamp(min=0)
        _close(y, ref)


def _ref_dgrad(dy_nhwc, cout, w_oihw, stride, pads, in_hw):
    n = dy_nhwc.shape[0]
    cin = w_oihw.shape[1]
    x = torch.zeros(n, cin, in_hw[0], in_hw[1], device="cuda", requires_grad=True)
    xp = F.pad(x, (pads[2], pads[3], pads[0], pads[1]))
    y = F.conv2d(xp, w_oihw.to(torch.bfloat16).float(), stride=stride)
    y.backward(dy_nhwc[..., :cout].float().permute(0, 3, 1, 2))
    return x.grad.permute(0, 2, 3, 1)


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride", [
    (2, 16, 16, 64, 256, 1, 1), (1, 25, 42, 256, 256, 3, 1), (2, 9, 10, 256, 36, 3, 1), (2, 7, 11, 256, 720, 3, 1),
    (2, 26, 42, 256, 128, 1, 2), (1, 25, 42, 128, 256, 3, 2), (1, 13, 21, 256, 256, 3, 2)])
def test_dgrad(n, h, w, cin, cout, k, stride):
    from batchai_retinanet_horovod_coco_b200.models.resnet import tf_same_pad
    wt = torch.randn(cout, cin, k, k, device="cuda") * (1.0 / (cout * k * k) ** 0.5)
    if stride == 1:
        pb, pads = (k // 2, k // 2), (k // 2,) * 4
        oh, ow = h, w
    else:
        ph, pw = tf_same_pad(h, k, 2), tf_same_pad(w, k, 2)
        pb, pads = (ph[0], pw[0]), (ph[0], ph[1], pw[0], pw[1])
        oh, ow = -(-h // 2), -(-w // 2)
    dy = _mk(n, oh, ow, cout)
    _, wd = cv.pack_working(cv.pack_master(wt))
    dx = cv.dgrad(dy, wd, cin, cout, k, k, (h, w), stride, pad=pb)
    ref = _ref_dgrad(dy, cout, wt, stride, pads, (h, w))
    _close(dx[..., :cin], ref)


def test_dgrad_mask_add():
    n, h, w, cin, cout = 2, 20, 30, 128, 64
    wt = torch.randn(cout, cin, 3, 3, device="cuda") * 0.05
    dy = _mk(n, h, w, cout)
    mask = _mk(n, h, w, cin, seed=5)
    add = _mk(n, h, w, cin, seed=6)
    _, wd = cv.pack_working(cv.pack_master(wt))
    dx = cv.dgrad(dy, wd, cin, cout, 3, 3, (h, w), 1, pad=(1, 1), mask=mask, add=add)
    ref = (_ref_dgrad(dy, cout, wt, 1, (1, 1, 1, 1), (h, w)) + add.float()) * (mask.float() > 0)
    _close(dx, ref)


def _ref_wgrad(x_nhwc, cin, dy_nhwc, cout, k, stride, pads):
    w = torch.zeros(cout, cin, k, k, device="cuda", requires_grad=True)
    x = F.pad(x_nhwc[..., :cin].float().permute(0, 3, 1, 2), (pads[2], pads[3], pads[0], pads[1]))
    y = F.conv2d(x, w, stride=stride)
    y.backward(dy_nhwc[..., :cout].float().permute(0, 3, 1, 2))
    return cv.pack_master(w.grad)


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride", [
    (2, 16, 16, 64, 64, 1, 1), (1, 25, 42, 256, 256, 3, 1), (2, 9, 10, 256, 36, 3, 1), (2, 7, 11, 256, 720, 3, 1),
    (2, 26, 42, 256, 128, 1, 2), (1, 25, 42, 128, 256, 3, 2), (2, 40, 56, 512, 128, 1, 1)])
def test_wgrad(n, h, w, cin, cout, k, stride):
    from batchai_retinanet_horovod_coco_b200.models.resnet import tf_same_pad
    x = _mk(n, h, w, cin)
    if stride == 1:
        pb, pads = (k // 2, k // 2), (k // 2,) * 4
        oh, ow = h, w
    else:
        ph, pw = tf_same_pad(h, k, 2), tf_same_pad(w, k, 2)
        pb, pads = (ph[0], pw[0]), (ph[0], ph[1], pw[0], pw[1])
        oh, ow = -(-h // 2), -(-w // 2)
    dy = _mk(n, oh, ow, cout, seed=2)
    rs = torch.rand(cout, device="cuda") + 0.5
    dw = torch.zeros(k * k, cout, cin, device="cuda")
    db = torch.zeros(cout, device="cuda")
    cv.wgrad(x, dy, dw, cin, cout, k, k, stride, pad=pb, rowscale=rs, dbias=db)
    ref = _ref_wgrad(x, cin, dy, cout, k, stride, pads) * rs.view(1, -1, 1)
    _close(dw, ref)
    _close(db, dy[..., :cout].float().sum(dim=(0, 1, 2)), tol=1e-3)


def test_wgrad_group():
    shapes = [(2, 25, 42), (2, 13, 21), (2, 7, 11), (2, 4, 6), (2, 2, 3)]
    xs = [_mk(n, h, w, 256, seed=i) for i, (n, h, w) in enumerate(shapes)]
    dys = [_mk(n, h, w, 256, seed=10 + i) for i, (n, h, w) in enumerate(shapes)]
    dw = torch.zeros(9, 256, 256, device="cuda")
    cv.wgrad_group(xs, dys, dw, 256, 256, 3, 3)
    ref = sum(_ref_wgrad(x, 256, dy, 256, 3, 1, (1, 1, 1, 1)) for x, dy in zip(xs, dys))
    _close(dw, ref)


def test_prepare_weights_kernel():
    m1 = torch.randn(9, 256, 64, device="cuda")
    m2 = torch.randn(1, 36, 256, device="cuda")
    s1 = torch.rand(256, device="cuda") + 0.5
    e = []
    outs = []
    for m, s in ((m1, s1), (m2, None)):
        t, co, ci = m.shape
        wf = torch.zeros(t, co, cv.round8(ci), device="cuda", dtype=torch.bfloat16)
        wd = torch.zeros(t, ci, cv.round8(co), device="cuda", dtype=torch.bfloat16)
        e.append(dict(src=m, scale=s, dst_f=wf, dst_d=wd, taps=t, Cout=co, Cin=ci))
        outs.append((wf, wd))
    tab = native.WeightTable(e)
    tab.run()
    for (m, s), (wf, wd) in zip(((m1, s1), (m2, None)), outs):
        rf, rd = cv.pack_working(m, s)
        assert torch.equal(wf, rf) and torch.equal(wd, rd)


@pytest.mark.parametrize("n,h,w", [(2, 64, 96), (1, 75, 101), (3, 256, 320)])
def test_stem_conv(n, h, w):
    torch.manual_seed(0)
    x = torch.zeros(n, h, w, 4, device="cuda", dtype=torch.bfloat16)
    x[..., :3] = (torch.randn(n, h, w, 3, device="cuda") * 60).to(torch.bfloat16)
    wt = torch.randn(64, 3, 7, 7, device="cuda") * 0.05
    scale = torch.rand(64, device="cuda") + 0.5
    shift = torch.randn(64, device="cuda")
    wp = native.pack_stem_weight(cv.pack_master(wt), scale)
    oh, ow = (h + 1) // 2, (w + 1) // 2
    y = torch.empty(n, oh, ow, 64, device="cuda", dtype=torch.bfloat16)
    native.stem_conv(x, wp, shift, y)
    w_eff = (wt * scale.view(64, 1, 1, 1)).to(torch.bfloat16).float()
    ref = F.conv2d(x[..., :3].float().permute(0, 3, 1, 2), w_eff, stride=2, padding=3).permute(0, 2, 3, 1) + shift
    _close(y, ref.clamp(min=0))


@pytest.mark.parametrize("n,h,w", [(2, 64, 96), (1, 75, 101), (3, 200, 264)])
def test_stem_wgrad(n, h, w):
    torch.manual_seed(0)
    x = torch.zeros(n, h, w, 4, device="cuda", dtype=torch.bfloat16)
    x[..., :3] = (torch.randn(n, h, w, 3, device="cuda") * 40).to(torch.bfloat16)
    oh, ow = (h + 1) // 2, (w + 1) // 2
    dy = (torch.randn(n, oh, ow, 64, device="cuda") * 0.1).to(torch.bfloat16)
    scale = torch.rand(64, device="cuda") + 0.5
    dw = torch.zeros(49, 64, 3, device="cuda")
    native.stem_wgrad(x, dy, scale, dw)
    wz = torch.zeros(64, 3, 7, 7, device="cuda", requires_grad=True)
    y = F.conv2d(x[..., :3].float().permute(0, 3, 1, 2), wz, stride=2, padding=3)
    y.backward(dy.float().permute(0, 3, 1, 2))
    ref = cv.pack_master(wz.grad) * scale.view(1, 64, 1)
    _close(dw, ref)


def _guarded(shape, dtype, fill=0.0):
    """A tensor carved out of the middle of a larger allocation whose guard zones hold a sentinel: kernels that
    write outside their output are caught without compute-sanitizer (which this GPU pool does not offer)."""
    n = 1
    for s in shape:
        n *= s
    guard = 4096
    raw = torch.full((n + 2 * guard,), 12345.0, device="cuda", dtype=dtype)
    view = raw[guard:guard + n].view(shape)
    view.fill_(fill)
    return raw, view, guard


def _guards_intact(raw, guard):
    return bool((raw[:guard] == 12345.0).all() and (raw[-guard:] == 12345.0).all())


def test_kernels_do_not_write_outside_their_outputs():
    """Odd sizes (partial tiles in every dimension), every kernel family, outputs surrounded by sentinels."""
    torch.manual_seed(3)
    n, h, w, cin, cout = 2, 13, 21, 128, 72
    x = _mk(n, h, w, cin)
    wt = torch.randn(cout, cin, 3, 3, device="cuda") * 0.03
    wf, wd = cv.pack_working(cv.pack_master(wt))
    # fprop stride 1 (+ shift, ReLU, residual)
    raw, y, g = _guarded((n, h, w, cv.round8(cout)), torch.bfloat16)
    res = _mk(n, h, w, cout, seed=5)
    native.conv_raw(0, [(x, y, res, None)], wf, cin, x.shape[-1], cout, y.shape[-1], 3, 3, 1, 1, 1,
                    shift=torch.zeros(cout, device="cuda"), relu=True)
    torch.cuda.synchronize()
    assert _guards_intact(raw, g)
    # dgrad stride 1 (+ mask)
    dy = _mk(n, h, w, cout, seed=6)
    raw, dx, g = _guarded((n, h, w, cin), torch.bfloat16)
    native.conv_raw(1, [(dy, dx, None, x)], wd, cout, dy.shape[-1], cin, dx.shape[-1], 3, 3, 1, 1, 1)
    torch.cuda.synchronize()
    assert _guards_intact(raw, g)
    # wgrad + bias gradient into fp32 buffers
    raw, dw, g = _guarded((9, cout, cin), torch.float32)
    rawb, db, gb = _guarded((cout,), torch.float32)
    native.wgrad_raw([(x, dy)], dw, cin, x.shape[-1], cout, dy.shape[-1], 3, 3, 1, 1, 1, dbias=db)
    torch.cuda.synchronize()
    assert _guards_intact(raw, g) and _guards_intact(rawb, gb) and dw.abs().sum() > 0
    # stem conv / stem wgrad at an odd size
    x4 = torch.zeros(1, 75, 101, 4, device="cuda", dtype=torch.bfloat16)
    x4[..., :3] = (torch.randn(1, 75, 101, 3, device="cuda") * 50).to(torch.bfloat16)
    wp = native.pack_stem_weight(torch.randn(49, 64, 3, device="cuda") * 0.05)
    raw, ys, g = _guarded((1, 38, 51, 64), torch.bfloat16)
    native.stem_conv(x4, wp, torch.zeros(64, device="cuda"), ys)
    raww, dws, gw = _guarded((49, 64, 3), torch.float32)
    native.stem_wgrad(x4, ys, torch.ones(64, device="cuda"), dws)
    torch.cuda.synchronize()
    assert _guards_intact(raw, g) and _guards_intact(raww, gw)
    # 2x2 pool forward / backward at odd sizes
    xp = torch.relu(_mk(2, 9, 13, 64, seed=7))
    raw, yp, g = _guarded((2, 4, 6, 64), torch.bfloat16)
    native.maxpool2_fwd(xp, yp)
    rawd, dxp, gd = _guarded((2, 9, 13, 64), torch.bfloat16)
    native.maxpool2_relu_bwd(xp, torch.ones_like(yp), dxp)
    torch.cuda.synchronize()
    assert _guards_intact(raw, g) and _guards_intact(rawd, gd)
