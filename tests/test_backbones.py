"""The non-ResNet backbones keras-retinanet offers through ``--backbone`` (reference train.py:323-324,390):
structure pinned against the published Keras parameter counts, pyramid taps, Keras ``.h5`` naming, and one
optimisation step through the CPU engine."""
import os

import pytest
import torch

from batchai_retinanet_horovod_coco_b200 import models
from batchai_retinanet_horovod_coco_b200.models.retinanet import make_backbone
from batchai_retinanet_horovod_coco_b200.utils import keras_h5

# keras.applications include_top=False totals (DenseNet without the trailing `bn`, which keras-retinanet's
# truncated graph drops: 4 * 1024 / 1664 / 1920 values)
KERAS_PARAMS = {
    "vgg16": 14_714_688,
    "vgg19": 20_024_384,
    "mobilenet224_1.0": 3_228_864,
    "densenet121": 7_037_504 - 4 * 1024,
    "densenet169": 12_642_880 - 4 * 1664,
    "densenet201": 18_321_984 - 4 * 1920,
}


@pytest.mark.parametrize("name", sorted(KERAS_PARAMS))
def test_parameter_counts_match_keras(name):
    m = make_backbone(name)
    n = sum(p.numel() for p in m.parameters()) + sum(b.numel() for b in m.buffers())
    assert n == KERAS_PARAMS[name]


@pytest.mark.parametrize("name,chans", [("vgg16", [256, 512, 512]), ("mobilenet160_0.5", [128, 256, 512]),
                                        ("densenet121", [512, 1024, 1024])])
def test_pyramid_taps(name, chans):
    m = make_backbone(name).eval()
    assert m.out_channels[1:] == chans
    for hw in ((97, 130), (64, 96)):
        with torch.no_grad():
            feats = m(torch.randn(1, 3, *hw))[1:]
        assert [f.shape[1] for f in feats] == chans
        assert [tuple(f.shape[-2:]) for f in feats] == m.feature_hw(*hw)


def test_keras_layer_names():
    assert list(make_backbone("vgg16").keras_layers())[:3] == ["block1_conv1", "block1_conv2", "block2_conv1"]
    mb = make_backbone("mobilenet224_0.25").keras_layers()
    assert {"conv1", "conv1_bn", "conv_dw_1", "conv_dw_13_bn", "conv_pw_13"} <= set(mb)
    dn = make_backbone("densenet121").keras_layers()
    assert {"conv1/conv", "conv1/bn", "conv2_block1_0_bn", "conv2_block6_2_conv", "pool2_conv", "pool4_bn",
            "conv5_block16_2_conv"} <= set(dn)


@pytest.mark.parametrize("name", ["vgg16", "mobilenet128_0.25", "densenet121"])
def test_h5_roundtrip_and_depthwise_layout(name, tmp_path):
    bb = models.backbone(name)
    m = bb.retinanet(3)
    path = os.path.join(tmp_path, "w.h5")
    keras_h5.save_weights(m, path)
    m2 = bb.retinanet(3)
    r = keras_h5.load_weights(m2, path, skip_mismatch=False)
    assert not r["missing"] and not r["skipped"]
    for a, b in zip(m.state_dict().values(), m2.state_dict().values()):
        assert torch.equal(a, b)
    if "mobilenet" in name:
        from batchai_retinanet_horovod_coco_b200.utils.h5 import H5File
        with H5File(path, "r") as f:
            k = f.dataset("/conv_dw_1/conv_dw_1/depthwise_kernel:0")
        assert k.shape == (3, 3, 8, 1)          # Keras DepthwiseConv2D layout (kh, kw, channels, multiplier)


@pytest.mark.parametrize("name", ["vgg16", "mobilenet128_0.25"])
def test_spec_trainer_step(name):
    from batchai_retinanet_horovod_coco_b200.engine import create_trainer
    torch.manual_seed(0)
    model = models.backbone(name).retinanet(4)
    tr = create_trainer(model, "cpu", lr=1e-4, clipnorm=0.0)
    img = torch.randint(0, 256, (1, 96, 128, 3), dtype=torch.uint8)
    gt = torch.tensor([[[10.0, 12.0, 70.0, 80.0, 1.0], [40.0, 30.0, 120.0, 90.0, 3.0]]])
    cnt = torch.tensor([2], dtype=torch.int32)
    hw = torch.tensor([[96, 128]], dtype=torch.int32)
    losses = [tr.train_step(img, gt, cnt, hw)[0].item() for _ in range(3)]
    assert all(l == l for l in losses) and losses[-1] < losses[0], losses
    b, s, l = tr.predict(img)
    assert b.shape == (1, 300, 4)


def test_trainable_bn_is_in_optimizer_weight_list():
    m = models.backbone("mobilenet128_0.25").retinanet(2)
    names = [n for n, _ in keras_h5.trainable_weight_list(m)]
    assert "conv1_bn/gamma:0" in names and "conv1_bn/beta:0" in names
    assert "conv1_bn/moving_mean:0" not in names
    r50 = models.backbone("resnet50").retinanet(2)
    assert not any("bn" in n for n, _ in keras_h5.trainable_weight_list(r50))   # frozen BN stays out
