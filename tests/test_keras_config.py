"""The Keras ``model_config`` written into full-model .h5 files (utils/keras_config.py), interpreted WITHOUT Keras: a
small shape-inference pass over the functional-Model JSON (the rules of each Keras layer class restated here) must
reproduce the PyTorch model -- same weighted layers, same kernel shapes given the inferred input channels (which fails
on any wrong connection), same output shapes."""
import json

import pytest
import torch

from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet
from batchai_retinanet_horovod_coco_b200.utils import keras_h5


def _conv_out(n, k, s, padding):
    return -(-n // s) if padding == "same" else (n - k) // s + 1


def interpret(cfg, input_shape, weights):
    """cfg: Model config; input_shape: (H, W, C).  Returns the shapes of the output tensors.  ``weights`` collects
    {layer name: expected Keras weight shapes} of every weighted layer met."""
    out = {}          # layer name -> list of node output shapes

    def get(ref):
        name, node, tensor = ref[0], ref[1], ref[2]
        assert name in out, f"layer {name} used before it is defined"
        assert tensor == 0 and node < len(out[name]), ref
        return out[name][node]

    assert len(cfg["input_layers"]) == 1
    for layer in cfg["layers"]:
        cls, c, name = layer["class_name"], layer["config"], layer["name"]
        assert c["name"] == name
        if cls == "InputLayer":
            assert c["batch_input_shape"] == [None, None, None, input_shape[2]] and layer["inbound_nodes"] == []
            out[name] = [tuple(input_shape)]
            continue
        shapes = []
        if cls == "Model":
            out[name] = [None]                   # node 0: the sub-model's own construction
            for node in layer["inbound_nodes"]:
                (src,) = node
                shapes.append(interpret(c, get(src), weights)[0])
            out[name] += shapes
            continue
        assert len(layer["inbound_nodes"]) == 1, name
        ins = [get(r) for r in layer["inbound_nodes"][0]]
        assert all(len(r) == 4 and r[3] == {} for r in layer["inbound_nodes"][0])
        x = ins[0]
        if cls in ("Conv2D", "DepthwiseConv2D"):
            (kh, kw), (sh, sw) = c["kernel_size"], c["strides"]
            assert c["data_format"] == "channels_last" and len(ins) == 1
            filters = x[2] if cls == "DepthwiseConv2D" else c["filters"]
            y = (_conv_out(x[0], kh, sh, c["padding"]), _conv_out(x[1], kw, sw, c["padding"]), filters)
            w = [(kh, kw, x[2], 1)] if cls == "DepthwiseConv2D" else [(kh, kw, x[2], filters)]
            if c["use_bias"]:
                w.append((filters,))
            weights[name] = w
        elif cls == "BatchNormalization":
            assert c["axis"] == 3
            y = x
            weights[name] = [(x[2],)] * 4
        elif cls == "Activation":
            assert c["activation"] in ("relu", "relu6", "sigmoid", "linear")
            y = x
        elif cls == "ZeroPadding2D":
            (t, b), (l, r) = c["padding"]
            y = (x[0] + t + b, x[1] + l + r, x[2])
        elif cls in ("MaxPooling2D", "AveragePooling2D"):
            (kh, kw), (sh, sw) = c["pool_size"], c["strides"]
            y = (_conv_out(x[0], kh, sh, c["padding"]), _conv_out(x[1], kw, sw, c["padding"]), x[2])
        elif cls == "Add":
            assert len(ins) == 2 and ins[0] == ins[1], (name, ins)
            y = x
        elif cls == "UpsampleLike":
            assert len(ins) == 2
            y = (ins[1][0], ins[1][1], x[2])
        elif cls == "Concatenate":
            ax = c["axis"] - 1           # batch axis removed
            assert all(s[:ax] + s[ax + 1:] == x[:ax] + x[ax + 1:] for s in ins), (name, ins)
            y = x[:ax] + (sum(s[ax] for s in ins),) + x[ax + 1:]
        elif cls == "Reshape":
            n = x[0] * x[1] * x[2]
            assert c["target_shape"][0] == -1 and n % c["target_shape"][1] == 0
            y = (n // c["target_shape"][1], c["target_shape"][1])
        else:
            raise AssertionError(f"unknown layer class {cls}")
        out[name] = [y]
    return [get(r) for r in cfg["output_layers"]]


@pytest.mark.parametrize("backbone", ["resnet50", "resnet101", "resnet152", "vgg16", "vgg19", "mobilenet224_1.0",
                                      "mobilenet128_0.25", "densenet121", "densenet169"])
def test_model_config_is_a_consistent_keras_graph(backbone):
    m = RetinaNet(7, backbone)
    cfg = json.loads(json.dumps(keras_h5.model_config(m)))          # through JSON, as it is stored
    assert cfg["class_name"] == "Model" and cfg["keras_version"] == "2.1.6" and cfg["backend"] == "tensorflow"
    c = cfg["config"]
    assert c["name"] == "retinanet" and c["input_layers"] == [["input_1", 0, 0]]
    assert c["output_layers"] == [["regression", 0, 0], ["classification", 0, 0]]
    H, W = 96, 128
    weights = {}
    reg, cls = interpret(c, (H, W, 3), weights)
    # same weighted layers as the module tree, and every weight shape follows from the GRAPH (input channels inferred)
    layers = m.keras_layers()
    assert set(weights) == set(layers)
    for name, shapes in weights.items():
        have = [tuple(a.shape) for _, a, _ in keras_h5._module_weights(name, layers[name])]
        assert have == shapes, (name, have, shapes)
    # output shapes: anchors x 4 and anchors x classes, as the PyTorch model produces them
    m.eval()
    with torch.no_grad():
        out = m(torch.zeros(1, 3, H, W))
    assert tuple(out["regression"].shape[1:]) == reg == (reg[0], 4)
    assert tuple(out["classification"].shape[1:]) == cls == (reg[0], 7)
    # keras-resnet's frozen batch norm and the head initialisers are there by name
    by_name = {l["name"]: l for l in c["layers"]}
    if backbone.startswith("resnet"):
        assert by_name["bn_conv1"]["config"]["freeze"] is True and by_name["bn_conv1"]["config"]["trainable"] is False
        assert by_name["bn_conv1"]["config"]["epsilon"] == pytest.approx(1e-5)
        assert by_name["padding_conv1"]["config"]["padding"] == [[3, 3], [3, 3]]
    sub = {l["name"]: l for l in by_name["classification_submodel"]["config"]["layers"]}
    assert sub["pyramid_classification"]["config"]["bias_initializer"] == {"class_name": "PriorProbability", "config": {"probability": 0.01}}
    assert len(by_name["classification_submodel"]["inbound_nodes"]) == 5
    assert [r[:3] for r in by_name["classification"]["inbound_nodes"][0]] == [["classification_submodel", i, 0] for i in range(1, 6)]


def test_full_model_file_carries_the_config(tmp_path):
    from batchai_retinanet_horovod_coco_b200.utils.h5 import H5File
    m = RetinaNet(3, "resnet50")
    p = str(tmp_path / "m.h5")
    keras_h5.save_model(m, p)
    with H5File(p, "r") as f:
        cfg = json.loads(bytes(f.attr("/", "model_config")).decode())
    assert len(cfg["config"]["layers"]) > 180
    raw = open(p, "rb").read()                  # ~100 KB of JSON: only possible as a variable-length string (global heap)
    at = raw.index(b"GCOL")
    assert raw[at + 4] == 1 and int.from_bytes(raw[at + 8:at + 16], "little") % 4096 == 0
    m2, _ = keras_h5.load_full_model(p)
    assert m2.backbone_name == "resnet50" and m2.num_classes == 3
