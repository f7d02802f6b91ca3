"""HDF5 subset + Keras checkpoint layout (round trips; byte-level structure checks from the format spec)."""
import json
import struct

import numpy as np
import pytest
import torch

from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.utils import keras_h5
from batchai_retinanet_horovod_coco_b200.utils.h5 import H5File


def test_raw_roundtrip_and_structure(tmp_path):
    p = str(tmp_path / "t.h5")
    a = np.arange(24, dtype=np.float32).reshape(2, 3, 4)
    with H5File(p, "w") as f:
        f.write_attr("/", "keras_version", b"2.1.6")
        f.write_attr("/", "layer_names", [b"conv1", b"bn_conv1_long_name"])
        f.write_attr("/", "vec", np.array([1.5, 2.5]))
        f.write_dataset("/conv1/conv1/kernel:0", a)
        f.write_dataset("/opt/iterations:0", np.asarray(7, dtype=np.int64))
        for i in range(70):                       # > one SNOD (32 entries): exercises the B-tree fan-out
            f.write_dataset(f"/many/d{i:03d}", np.full((3,), i, dtype=np.float64))
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and raw[8] == 0            # superblock v0
    assert raw[13] == 8 and raw[14] == 8
    eof = struct.unpack_from("<Q", raw, 40)[0]
    assert eof == len(raw)
    root_oh = struct.unpack_from("<Q", raw, 64)[0]
    assert raw[root_oh] == 1                                           # v1 object header
    bt, heap = struct.unpack_from("<QQ", raw, 80)
    assert raw[bt:bt + 4] == b"TREE" and raw[heap:heap + 4] == b"HEAP"
    with H5File(p, "r") as f:
        assert sorted(f.keys("/")) == ["conv1", "many", "opt"]
        assert f.attr("/", "keras_version") == b"2.1.6"
        assert [bytes(x) for x in f.attr("/", "layer_names")] == [b"conv1", b"bn_conv1_long_name"]
        assert np.allclose(f.attr("/", "vec"), [1.5, 2.5])
        assert np.array_equal(f.dataset("/conv1/conv1/kernel:0"), a)
        assert int(f.dataset("/opt/iterations:0")) == 7
        assert len(f.keys("/many")) == 70 and f.keys("/many")[0] == "d000"
        assert np.array_equal(f.dataset("/many/d069"), np.full((3,), 69.0))
        assert f.is_group("/conv1/conv1") and not f.is_group("/conv1/conv1/kernel:0")
        with pytest.raises(KeyError):
            f.dataset("/nope")


def test_rejects_non_hdf5(tmp_path):
    p = tmp_path / "x.h5"
    p.write_bytes(b"not hdf5" * 100)
    with pytest.raises(OSError):
        H5File(str(p), "r")


@pytest.fixture(scope="module")
def model():
    torch.manual_seed(0)
    m = RetinaNet(80, "resnet50")
    with torch.no_grad():
        for b in m.modules():
            if hasattr(b, "moving_variance"):
                b.moving_mean.normal_(0, 0.1)
                b.moving_variance.uniform_(0.5, 1.5)
    return m


def test_weights_roundtrip_by_name(tmp_path, model):
    p = str(tmp_path / "w.h5")
    keras_h5.save_weights(model, p)
    with H5File(p, "r") as f:
        names = [bytes(x).decode() for x in f.attr("/", "layer_names")]
        assert names[0] == "conv1" and "regression_submodel" in names and "classification_submodel" in names
        k = f.dataset("/conv1/conv1/kernel:0")
        assert k.shape == (7, 7, 3, 64)                                     # HWIO
        wn = [bytes(x).decode() for x in f.attr("/classification_submodel", "weight_names")]
        assert wn[0] == "pyramid_classification_0/kernel:0" and wn[-1] == "pyramid_classification/bias:0"
    m2 = RetinaNet(80, "resnet50")
    rep = keras_h5.load_weights(m2, p)
    assert not rep["skipped"] and not rep["missing"]
    for (n1, t1), (n2, t2) in zip(model.state_dict().items(), m2.state_dict().items()):
        assert n1 == n2 and torch.equal(t1, t2), n1


def test_skip_mismatch_for_other_class_count(tmp_path, model):
    p = str(tmp_path / "w.h5")
    keras_h5.save_weights(model, p)
    m20 = RetinaNet(20, "resnet50")
    with pytest.warns(UserWarning):
        rep = keras_h5.load_weights(m20, p, skip_mismatch=True)
    assert sorted(rep["skipped"]) == ["pyramid_classification/bias:0", "pyramid_classification/kernel:0"]
    assert torch.equal(m20.backbone.conv1.weight, model.backbone.conv1.weight)
    with pytest.raises(ValueError):
        keras_h5.load_weights(RetinaNet(20, "resnet50"), p, skip_mismatch=False)


def test_full_model_with_optimizer_state(tmp_path, model):
    p = str(tmp_path / "checkpoint-01.h5")
    tw = keras_h5.trainable_weight_list(model)
    assert len(tw) == 89
    slots = {n: (torch.full_like(t, 0.25), torch.full_like(t, 0.5)) for n, t in tw[:3]}
    cfg = {"lr": 1e-5, "beta_1": 0.9, "beta_2": 0.999, "epsilon": 1e-7, "decay": 0.0, "clipnorm": 0.001}
    keras_h5.save_model(model, p, optimizer_config=cfg, optimizer_slots=slots, iterations=500, epoch=1)
    m2, st = keras_h5.load_full_model(p)
    assert m2.num_classes == 80 and m2.backbone_name == "resnet50"
    assert st["iterations"] == 500 and st["epoch"] == 1 and st["optimizer_config"]["clipnorm"] == 0.001
    assert len(st["slots"]) == 89
    m, v = st["slots"][tw[0][0]]
    assert m.shape == tw[0][1].shape and float(m.mean()) == 0.25 and float(v.mean()) == 0.5
    assert float(st["slots"][tw[5][0]][0].abs().sum()) == 0
    assert torch.equal(m2.fpn.P6.weight, model.fpn.P6.weight)
    with H5File(p, "r") as f:
        tc = json.loads(bytes(f.attr("/", "training_config")).decode())
        assert tc["loss"] == {"regression": "_smooth_l1", "classification": "_focal"}
        assert bytes(f.attr("/optimizer_weights", "weight_names")[0]) == b"Adam/iterations:0"
