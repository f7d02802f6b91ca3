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
        assert int(np.asarray(f.dataset("/opt/iterations:0")).reshape(-1)[0]) == 7
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


# ---------------------------------------------------------------------------------------------------------------------
# A file this package's writer did NOT produce: tests/fixtures/keras_resnet50_stem_libhdf5.h5 is assembled byte by byte
# from the HDF5 format specification in libhdf5 1.8's layout (tests/fixtures/libhdf5_layout.py lists what that means:
# header continuation blocks, NIL messages, 8-byte padded v1 attributes, fixed-length string arrays, dataspaces with
# maximum dimensions, fill-value / mtime messages, several symbol-table leaves, cached group entries).
def _fixture():
    import os, sys
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fixtures")
    sys.path.insert(0, d)
    import make_keras_fixture as mk
    return mk


def test_foreign_layout_fixture_structure():
    mk = _fixture()
    raw = open(mk.PATH, "rb").read()
    assert raw == mk.build(), "committed fixture is out of date: python tests/fixtures/make_keras_fixture.py"
    # the features the fixture exists for are really in the bytes
    root_oh = struct.unpack_from("<Q", raw, 64)[0]
    assert root_oh == 96 and raw[root_oh] == 1
    nmsg, _, size = struct.unpack_from("<HII", raw, root_oh + 2)
    types, pos = [], root_oh + 16
    while pos < root_oh + 16 + size:
        t, n = struct.unpack_from("<HH", raw, pos)
        types.append(t)
        if t == 0x10:
            cont, clen = struct.unpack_from("<QQ", raw, pos + 8)
        pos += 8 + n
    assert types == [0x11, 0x00, 0x10], types                 # symbol table, NIL filler, continuation
    assert cont > root_oh + 16 + size and cont + clen <= len(raw)
    t0, n0 = struct.unpack_from("<HH", raw, cont)
    assert t0 == 0x0C                                         # the attributes live in the continuation block
    name_size, = struct.unpack_from("<H", raw, cont + 8 + 2)
    assert raw[cont + 16:cont + 16 + 12] == b"layer_names\0" and name_size == 12      # 12 bytes, padded to 16
    assert nmsg == 3 + 3 + 1
    at = raw.index(b"GCOL")                                   # backend / keras_version: variable-length strings
    assert struct.unpack_from("<Q", raw, at + 8)[0] == 4096 and raw[at + 16 + 16:at + 16 + 26] == b"tensorflow"
    with H5File(mk.PATH, "r") as f:
        assert f.keys("/") == sorted(n for n, _ in mk.LAYERS)
        assert [bytes(x).decode() for x in f.attr("/", "layer_names")] == [n for n, _ in mk.LAYERS]
        assert f.attr("/", "backend") == b"tensorflow" and f.attr("/", "keras_version") == b"2.1.6"
        assert len(f.attr("/pool1", "weight_names")) == 0
        for layer, weights in mk.LAYERS:
            assert [bytes(x).decode() for x in np.atleast_1d(f.attr("/" + layer, "weight_names"))] == [w for w, _ in weights]
            for w, shape in weights:
                assert np.array_equal(f.dataset(f"/{layer}/{w}"), mk.values(w, shape))


def test_foreign_layout_fixture_loads_by_name():
    mk = _fixture()
    torch.manual_seed(1)
    m = RetinaNet(80, "resnet50")
    before = m.keras_layers()["res2a_branch2b"].weight.detach().clone()
    with pytest.warns(UserWarning, match="mismatch in shape"):
        rep = keras_h5.load_weights(m, mk.PATH, by_name=True, skip_mismatch=True)
    assert set(rep["loaded"]) == {"conv1/kernel:0", "res2a_branch2a/kernel:0"} | {
        f"{bn}/{k}:0" for bn in ("bn_conv1", "bn2a_branch2a") for k in ("gamma", "beta", "moving_mean", "moving_variance")}
    assert rep["skipped"] == ["res2a_branch2b/kernel:0"]      # (3, 3, 64, 32) in the file, (3, 3, 64, 64) in the model
    assert "res2b_branch2a/kernel:0" in rep["missing"] and not any(w.startswith("fc1000") for w in rep["missing"])
    layers = m.keras_layers()
    k = mk.values("conv1/kernel:0", (7, 7, 3, 64))             # Keras HWIO -> OIHW
    assert torch.equal(layers["conv1"].weight.detach(), torch.from_numpy(k.transpose(3, 2, 0, 1).copy()))
    assert torch.equal(layers["bn_conv1"].moving_variance.detach(),
                       torch.from_numpy(mk.values("bn_conv1/moving_variance:0", (64,))))
    assert torch.equal(layers["res2a_branch2b"].weight.detach(), before)


def test_reader_survives_damaged_files():
    """300 truncated / bit-flipped copies of the foreign-layout fixture: a clean read or a Python exception each time,
    never a crash (shape fields are validated against the stored data, ranks and sizes bounded)."""
    import os, subprocess, sys
    w = os.path.join(os.path.dirname(os.path.abspath(__file__)), "workers", "h5_fuzz.py")
    r = subprocess.run([sys.executable, w, "300"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "FUZZ_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    raised = int(r.stdout.split("raised=")[1].split()[0])
    assert raised > 100          # the damage is real: most cases must be rejected, not silently read
