"""Spec layer (plain PyTorch, CPU): structure, anchors, targets, losses, detections.
Hand-computed fixtures in the style of keras-retinanet's own tests."""
import math

import pytest
import torch

from batchai_retinanet_horovod_coco_b200.models import backbone
from batchai_retinanet_horovod_coco_b200.models.resnet import block_name, tf_same_pad
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet, retinanet_bbox, retinanet_forward_macs
from batchai_retinanet_horovod_coco_b200.ops import anchors as A
from batchai_retinanet_horovod_coco_b200.ops import losses as L
from batchai_retinanet_horovod_coco_b200.ops.boxes import bbox_transform_inv, clip_boxes
from batchai_retinanet_horovod_coco_b200.ops.filter_detections import filter_detections, nms_reference


@pytest.fixture(scope="module")
def r50():
    torch.manual_seed(0)
    return RetinaNet(80, "resnet50")


def test_parameter_inventory_matches_survey(r50):
    trainable = sum(p.numel() for p in r50.parameters())
    assert trainable == 37_915_572
    assert len(list(r50.parameters())) == 89
    assert sum(b.numel() for b in r50.buffers()) == 106_240  # frozen BN vectors
    names = list(r50.keras_layers())
    assert names[:4] == ["conv1", "bn_conv1", "res2a_branch2a", "bn2a_branch2a"]
    assert "res5c_branch2c" in names and "P7" in names and "pyramid_classification" in names


def test_block_names_long_stages():
    assert block_name(0, 2, 3) == "2c"
    assert block_name(2, 0, 23) == "4a" and block_name(2, 5, 23) == "4b5"


def test_forward_macs_match_survey():
    m = retinanet_forward_macs("resnet50", 800, 1333)
    assert abs(m["total"] / 1e9 - 244.6) < 0.1 and abs(m["heads"] / 1e9 - 144.1) < 0.1


def test_tf_same_pad():
    assert tf_same_pad(400, 3, 2) == (0, 1)
    assert tf_same_pad(667, 3, 2) == (1, 1)
    assert tf_same_pad(42, 3, 2) == (0, 1) and tf_same_pad(25, 3, 2) == (1, 1)


def test_output_shapes_and_prior_bias(r50):
    x = torch.randn(1, 3, 200, 264)
    out = r50(x)
    shapes = A.guess_shapes((200, 264))
    assert out["feature_shapes"] == [tuple(s) for s in shapes]
    n = sum(h * w * 9 for h, w in shapes)
    assert out["regression"].shape == (1, n, 4) and out["classification"].shape == (1, n, 80)
    b = r50.classification_submodel.out.bias
    assert torch.allclose(b, torch.full_like(b, -math.log(99.0)))


def test_freeze_backbone(r50):
    m = RetinaNet(3, "resnet50").freeze_backbone()
    t = m.trainable_parameters()
    assert len(t) == 36
    assert sum(p.numel() for p in t) == 14_460_660 - (9 * 80 - 9 * 3) * (256 * 9 + 1)


def test_backbone_factory():
    assert backbone("resnet101").retinanet(2).backbone_name == "resnet101"
    with pytest.raises(NotImplementedError):
        backbone("inception")
    with pytest.raises(ValueError):
        backbone("resnet18")
    with pytest.raises(ValueError):
        backbone("vgg11")
    with pytest.raises(ValueError):
        backbone("mobilenet224_0.3")


def test_generate_anchors_fixture():
    a = A.generate_anchors(32)
    assert a.shape == (9, 4)
    # ratio 0.5, scale 1: area 1024 -> w = sqrt(2048), h = w/2
    w = math.sqrt(2048.0)
    assert torch.allclose(a[0], torch.tensor([-w / 2, -w / 4, w / 2, w / 4], dtype=torch.float64))
    # ratio 1, scale 2^(1/3)
    s = 32 * 2 ** (1 / 3)
    assert torch.allclose(a[4], torch.tensor([-s / 2, -s / 2, s / 2, s / 2], dtype=torch.float64))


def test_anchor_count_and_order():
    assert A.anchors_for_shape((800, 1333)).shape == (200700, 4)
    a = A.anchors_for_shape((64, 64))
    # level P3 is 8x8: second pixel's first anchor is the first anchor shifted by the stride in x
    assert torch.allclose(a[9] - a[0], torch.tensor([8.0, 0.0, 8.0, 0.0]))
    assert torch.allclose(a[8 * 9] - a[0], torch.tensor([0.0, 8.0, 0.0, 8.0]))


def test_compute_overlap_plus_one():
    b = torch.tensor([[0.0, 0.0, 9.0, 9.0]])
    q = torch.tensor([[0.0, 0.0, 9.0, 9.0], [5.0, 5.0, 14.0, 14.0], [20.0, 20.0, 30.0, 30.0]])
    iou = A.compute_overlap(b, q)
    assert abs(iou[0, 0].item() - 1.0) < 1e-6
    assert abs(iou[0, 1].item() - 25.0 / 175.0) < 1e-6   # +1 convention: 5x5 / (100+100-25)
    assert iou[0, 2].item() == 0.0


def test_assignment_rules_and_targets():
    anchors = A.anchors_for_shape((128, 128))
    gt = torch.tensor([[20.0, 30.0, 90.0, 100.0, 5.0]])
    state, matched, iou = A.assign_targets(anchors, gt, (128, 100))
    assert set(state.unique().tolist()) <= {-1, 0, 1}
    assert ((iou >= 0.5) & (state == 0)).sum() == 0
    cx = (anchors[:, 0] + anchors[:, 2]) / 2
    assert (state[cx >= 100] == -1).all()                 # centre outside the un-padded image -> ignore
    reg, lab = A.anchor_targets_bbox(anchors, [gt], 80, [(128, 100)])
    pos = lab[0, :, -1] == 1
    assert pos.any() and (lab[0, pos, 5] == 1).all() and lab[0, pos, :80].sum() == pos.sum()
    a = anchors[pos][0]
    t = reg[0, pos][0]
    assert abs(t[0].item() - (20.0 - a[0]) / (a[2] - a[0]) / 0.2) < 1e-4
    # empty annotations: everything background
    s0, _, _ = A.assign_targets(anchors, torch.zeros(0, 5), None)
    assert (s0 == 0).all()


def test_focal_dense_and_compact_agree():
    torch.manual_seed(0)
    B, An, K = 2, 50, 4
    logits = torch.randn(B, An, K)
    state = torch.randint(-1, 2, (B, An)).to(torch.int8)
    cls = torch.randint(0, K, (B, An))
    y_true = torch.zeros(B, An, K + 1)
    y_true[..., -1] = state.float()
    pos = state == 1
    y_true[..., :K][pos] = torch.nn.functional.one_hot(cls[pos], K).float()
    dense = L.focal(y_true, torch.sigmoid(logits))
    compact = L.focal_compact(logits, state, cls)
    assert abs(dense.item() - compact.item()) < 1e-5 * max(1.0, abs(dense.item()))


def test_focal_hand_value():
    # one positive anchor with p = 0.5 on its class, one background class with p = 0.5
    y_true = torch.tensor([[[1.0, 0.0, 1.0]]])
    y_pred = torch.tensor([[[0.5, 0.5]]])
    want = 0.25 * 0.25 * math.log(2) + 0.75 * 0.25 * math.log(2)
    assert abs(L.focal(y_true, y_pred).item() - want) < 1e-6


def test_smooth_l1_hand_value():
    y_true = torch.tensor([[[0.0, 0.0, 0.0, 0.0, 1.0], [0.0, 0.0, 0.0, 0.0, 0.0]]])
    y_pred = torch.tensor([[[0.05, 1.0, -2.0, 0.0], [9.0, 9.0, 9.0, 9.0]]])
    want = 0.5 * 9 * 0.05 ** 2 + (1.0 - 0.5 / 9) + (2.0 - 0.5 / 9)
    assert abs(L.smooth_l1(y_true, y_pred).item() - want) < 1e-6
    t = y_true[..., :4]
    assert abs(L.smooth_l1_compact(y_pred, t, y_true[..., 4]).item() - want) < 1e-6


def test_box_decode_roundtrip():
    anchors = A.anchors_for_shape((64, 64))[:20]
    gt = anchors + torch.tensor([3.0, -2.0, 5.0, 4.0])
    t = A.bbox_transform(anchors, gt)
    assert torch.allclose(bbox_transform_inv(anchors, t), gt, atol=1e-4)
    c = clip_boxes(torch.tensor([[-5.0, -5.0, 700.0, 700.0]]), (64, 32))
    assert c.tolist() == [[0.0, 0.0, 32.0, 64.0]]


def test_nms_and_filter_detections():
    boxes = torch.tensor([[0.0, 0, 10, 10], [1, 1, 11, 11], [20, 20, 30, 30], [0, 0, 10, 10]])
    scores = torch.tensor([0.9, 0.8, 0.7, 0.6])
    assert nms_reference(boxes, scores, 0.5).tolist() == [0, 2]
    cls = torch.zeros(1, 4, 2)
    cls[0, :, 0] = scores
    cls[0, 3, 1] = 0.95                                   # same box, other class: survives class-specific NMS
    b, s, l = filter_detections(boxes.unsqueeze(0), cls)
    assert b.shape == (1, 300, 4) and s.shape == (1, 300) and l.shape == (1, 300)
    k = int((l[0] >= 0).sum())
    assert k == 3 and l[0, :3].tolist() == [1, 0, 0] and abs(s[0, 0].item() - 0.95) < 1e-6
    assert (l[0, 3:] == -1).all() and (b[0, 3:] == -1).all()


def test_prediction_model_runs(r50):
    pm = retinanet_bbox(r50)
    b, s, l = pm(torch.randn(1, 3, 128, 160))
    assert b.shape == (1, 300, 4) and s.shape == (1, 300) and l.shape == (1, 300)


def test_affine_inversion_matches_opencv():
    """``ops.native.invert_affine_like_cv2`` (host side of the GPU warp: the dst -> src map) against OpenCV's own
    inversion, for the transforms the augmentation generates and for a singular matrix."""
    import cv2
    import numpy as np
    from batchai_retinanet_horovod_coco_b200.ops.native import invert_affine_like_cv2
    from batchai_retinanet_horovod_coco_b200.utils.transform import random_transform_generator
    gen = random_transform_generator(min_rotation=-0.1, max_rotation=0.1, min_translation=(-0.1, -0.1),
                                     max_translation=(0.1, 0.1), min_shear=-0.1, max_shear=0.1, min_scaling=(0.9, 0.9),
                                     max_scaling=(1.1, 1.1), flip_x_chance=0.5, flip_y_chance=0.5,
                                     prng=np.random.RandomState(3))
    for _ in range(50):
        m = np.asarray(next(gen), dtype=np.float64)[:2, :]
        assert np.allclose(invert_affine_like_cv2(m), cv2.invertAffineTransform(m), rtol=1e-12, atol=1e-12)
    assert np.array_equal(invert_affine_like_cv2([[1, 2, 3], [2, 4, 5]])[:, :2], np.zeros((2, 2)))     # singular: D = 0


def test_image_geometry_restatement_matches_opencv():
    """``ops.image_ref`` (integer-for-integer NumPy restatement of csrc/image_sm100.cu) against OpenCV over many random
    sizes, scales and affine maps: the warp is bit-exact, the resize differs by at most 1 LSB on < 0.5 % of the values
    (OpenCV's vectorised path rounds a few half-way cases differently)."""
    import cv2
    import numpy as np
    from batchai_retinanet_horovod_coco_b200.ops import image_ref
    from batchai_retinanet_horovod_coco_b200.ops.native import invert_affine_like_cv2
    rng = np.random.default_rng(0)
    bad = tot = 0
    for t in range(60):
        h, w = (int(v) for v in rng.integers(20, 260, 2))
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        if t % 2:
            img = cv2.GaussianBlur(img, (5, 5), 1.2)
        sc = float(rng.uniform(0.4, 3.0))
        want = cv2.resize(img, None, fx=sc, fy=sc)
        got = image_ref.resize_linear_u8(img, want.shape[0], want.shape[1], 1.0 / sc, 1.0 / sc)
        d = np.abs(got.astype(int) - want.astype(int))
        assert d.max() <= 1
        bad, tot = bad + int((d > 0).sum()), tot + d.size
    assert bad / tot < 5e-3
    for t in range(60):
        h, w = (int(v) for v in rng.integers(20, 260, 2))
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        a, s, k = rng.uniform(-0.3, 0.3), rng.uniform(0.8, 1.25), rng.uniform(-0.2, 0.2)
        M = np.array([[s * np.cos(a), -s * np.sin(a + k), rng.uniform(-0.2, 0.2) * w],
                      [s * np.sin(a), s * np.cos(a + k), rng.uniform(-0.2, 0.2) * h]])
        if t % 3 == 0:                                   # horizontal flip
            M[0] *= -1
            M[0, 2] += w
        for rep in (True, False):
            want = cv2.warpAffine(img, M, dsize=(w, h), flags=cv2.INTER_LINEAR,
                                  borderMode=cv2.BORDER_REPLICATE if rep else cv2.BORDER_CONSTANT, borderValue=(7, 7, 7))
            got = image_ref.warp_affine_u8(img, invert_affine_like_cv2(M), rep, 7)
            assert np.array_equal(got, want), (t, rep)
