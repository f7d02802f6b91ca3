"""Keras 2.1.6 functional-``Model`` configuration of a RetinaNet, as ``model.to_json()`` / ``ModelCheckpoint`` store it in
the ``model_config`` attribute of a full-model ``.h5`` (reference ``train.py:113-114`` writes such files, ``train.py:406``
reopens them with ``keras_retinanet.models.load_model``).

The graph is re-derived layer by layer from the module tree: every Keras layer of keras-resnet / keras.applications /
keras-retinanet appears with its class name, its full constructor configuration and its ``inbound_nodes``, including the
layers without weights (``ZeroPadding2D``, ``Activation``, ``Add``, ``MaxPooling2D``, ``UpsampleLike``, ``Reshape``,
``Concatenate``) and the two head sub-``Model``s applied to the five pyramid levels (node indices 1..5).  Custom classes
(``BatchNormalization`` of keras-resnet with ``freeze``, ``UpsampleLike``, ``PriorProbability``, ``relu6`` /
``DepthwiseConv2D`` of keras.applications.mobilenet) are referenced by the names keras-retinanet registers in
``custom_objects``.

``tests/test_h5.py`` interprets the emitted JSON with an independent shape-inference pass (no Keras in this image) and
checks layer order, connectivity, weight shapes and the output shapes against the PyTorch model.
"""
from __future__ import annotations

import torch.nn as nn

_ZEROS = {"class_name": "Zeros", "config": {}}
_ONES = {"class_name": "Ones", "config": {}}
_GLOROT = {"class_name": "VarianceScaling", "config": {"scale": 1.0, "mode": "fan_avg", "distribution": "uniform", "seed": None}}
_NORMAL001 = {"class_name": "RandomNormal", "config": {"mean": 0.0, "stddev": 0.01, "seed": None}}


def _trainable(mod) -> bool:
    ps = list(mod.parameters(recurse=False))
    return bool(ps) and all(p.requires_grad for p in ps)


class Graph:
    """Records layers the way Keras' functional API does: ``add`` returns the reference of the new layer's output."""

    def __init__(self, name: str):
        self.name = name
        self.layers = []
        self.index = {}
        self.inputs = []

    def add(self, class_name: str, name: str, config: dict, inbound=None, trainable=True):
        cfg = {"name": name, "trainable": bool(trainable)}
        cfg.update(config)
        nodes = []
        if inbound is not None:
            refs = inbound if isinstance(inbound, list) else [inbound]
            nodes = [[[r[0], r[1], r[2], {}] for r in refs]]
        assert name not in self.index, f"duplicate layer name {name}"
        self.index[name] = len(self.layers)
        self.layers.append({"name": name, "class_name": class_name, "config": cfg, "inbound_nodes": nodes})
        return (name, 0, 0)

    def input(self, name: str, shape):
        self.layers.append({"name": name, "class_name": "InputLayer",
                            "config": {"batch_input_shape": [None] + list(shape), "dtype": "float32", "sparse": False, "name": name},
                            "inbound_nodes": []})
        self.index[name] = len(self.layers) - 1
        self.inputs.append([name, 0, 0])
        return (name, 0, 0)

    def config(self, outputs) -> dict:
        return {"name": self.name, "layers": self.layers, "input_layers": self.inputs,
                "output_layers": [[o[0], o[1], o[2]] for o in outputs]}

    # ------------------------------------------------------------------ layer helpers
    def conv(self, name, mod: nn.Conv2d, x, padding="valid", activation="linear", kernel_init=None, bias_init=None):
        cfg = {"filters": mod.weight.shape[0], "kernel_size": list(mod.kernel_size), "strides": list(mod.stride),
               "padding": padding, "data_format": "channels_last", "dilation_rate": [1, 1], "activation": activation,
               "use_bias": mod.bias is not None, "kernel_initializer": kernel_init or _GLOROT,
               "bias_initializer": bias_init or _ZEROS, "kernel_regularizer": None, "bias_regularizer": None,
               "activity_regularizer": None, "kernel_constraint": None, "bias_constraint": None}
        return self.add("Conv2D", name, cfg, x, _trainable(mod))

    def depthwise(self, name, mod: nn.Conv2d, x):
        cfg = {"kernel_size": list(mod.kernel_size), "strides": list(mod.stride), "padding": "same",
               "data_format": "channels_last", "activation": "linear", "use_bias": False, "depth_multiplier": 1,
               "depthwise_initializer": _GLOROT, "bias_initializer": _ZEROS, "depthwise_regularizer": None,
               "bias_regularizer": None, "activity_regularizer": None, "depthwise_constraint": None, "bias_constraint": None}
        return self.add("DepthwiseConv2D", name, cfg, x, _trainable(mod))

    def bn(self, name, mod, x, freeze=None):
        cfg = {"axis": 3, "momentum": float(getattr(mod, "momentum", 0.99)), "epsilon": float(mod.eps), "center": True,
               "scale": True, "beta_initializer": _ZEROS, "gamma_initializer": _ONES, "moving_mean_initializer": _ZEROS,
               "moving_variance_initializer": _ONES, "beta_regularizer": None, "gamma_regularizer": None,
               "beta_constraint": None, "gamma_constraint": None}
        trainable = _trainable(mod)
        if freeze is not None:                      # keras_resnet.layers.BatchNormalization
            cfg["freeze"] = bool(freeze)
            trainable = not freeze
        return self.add("BatchNormalization", name, cfg, x, trainable)

    def act(self, name, fn, x):
        return self.add("Activation", name, {"activation": fn}, x)

    def pad(self, name, p, x):
        return self.add("ZeroPadding2D", name, {"padding": [[p, p], [p, p]], "data_format": "channels_last"}, x)

    def pool(self, cls, name, size, stride, padding, x):
        return self.add(cls, name, {"pool_size": [size, size], "padding": padding, "strides": [stride, stride],
                                    "data_format": "channels_last"}, x)


# ---------------------------------------------------------------------------------------------------------------------
def _resnet(g: Graph, bb, x):
    x = g.pad("padding_conv1", 3, x)
    x = g.conv("conv1", bb.conv1, x)
    x = g.bn("bn_conv1", bb.bn_conv1, x, freeze=True)
    x = g.act("conv1_relu", "relu", x)
    x = g.pool("MaxPooling2D", "pool1", 3, 2, "same", x)
    outs = []
    for stage in bb.stages:
        for blk in stage:
            n = blk.name
            y = g.conv(f"res{n}_branch2a", blk.branch2a, x)
            y = g.bn(f"bn{n}_branch2a", blk.bn2a, y, freeze=True)
            y = g.act(f"res{n}_branch2a_relu", "relu", y)
            y = g.pad(f"padding{n}_branch2b", 1, y)
            y = g.conv(f"res{n}_branch2b", blk.branch2b, y)
            y = g.bn(f"bn{n}_branch2b", blk.bn2b, y, freeze=True)
            y = g.act(f"res{n}_branch2b_relu", "relu", y)
            y = g.conv(f"res{n}_branch2c", blk.branch2c, y)
            y = g.bn(f"bn{n}_branch2c", blk.bn2c, y, freeze=True)
            sc = x
            if blk.project:
                sc = g.conv(f"res{n}_branch1", blk.branch1, x)
                sc = g.bn(f"bn{n}_branch1", blk.bn1, sc, freeze=True)
            x = g.add("Add", f"res{n}", {}, [y, sc])
            x = g.act(f"res{n}_relu", "relu", x)
        outs.append(x)
    return outs[1:]


def _vgg(g: Graph, bb, x):
    names = iter(bb._names)
    outs = []
    for b, convs in enumerate(bb.blocks, start=1):
        for c in convs:
            x = g.conv(next(names), c, x, padding="same", activation="relu")
        x = g.pool("MaxPooling2D", f"block{b}_pool", 2, 2, "valid", x)
        outs.append(x)
    return outs[2:]


def _mobilenet(g: Graph, bb, x):
    x = g.conv("conv1", bb.conv1, x, padding="same")
    x = g.bn("conv1_bn", bb.conv1_bn, x)
    x = g.act("conv1_relu", "relu6", x)
    taps = {}
    for i in range(13):
        x = g.depthwise(f"conv_dw_{i + 1}", bb.dw[i], x)
        x = g.bn(f"conv_dw_{i + 1}_bn", bb.dw_bn[i], x)
        x = g.act(f"conv_dw_{i + 1}_relu", "relu6", x)
        x = g.conv(f"conv_pw_{i + 1}", bb.pw[i], x, padding="same")
        x = g.bn(f"conv_pw_{i + 1}_bn", bb.pw_bn[i], x)
        x = g.act(f"conv_pw_{i + 1}_relu", "relu6", x)
        taps[i + 1] = x
    return [taps[5], taps[11], taps[13]]


def _densenet(g: Graph, bb, x):
    x = g.pad("zero_padding2d_1", 3, x)
    x = g.conv("conv1/conv", bb.conv1, x)
    x = g.bn("conv1/bn", bb.bn1, x)
    x = g.act("conv1/relu", "relu", x)
    x = g.pad("zero_padding2d_2", 1, x)
    x = g.pool("MaxPooling2D", "pool1", 3, 2, "valid", x)
    outs = []
    for bi, layers in enumerate(bb.dense):
        for li, l in enumerate(layers):
            p = f"conv{bi + 2}_block{li + 1}"
            y = g.bn(f"{p}_0_bn", l.bn0, x)
            y = g.act(f"{p}_0_relu", "relu", y)
            y = g.conv(f"{p}_1_conv", l.conv1, y)
            y = g.bn(f"{p}_1_bn", l.bn1, y)
            y = g.act(f"{p}_1_relu", "relu", y)
            y = g.conv(f"{p}_2_conv", l.conv2, y, padding="same")
            x = g.add("Concatenate", f"{p}_concat", {"axis": 3}, [x, y])
        outs.append(x)
        if bi < 3:
            t = bb.trans[bi]
            x = g.bn(f"pool{bi + 2}_bn", t.bn, x)
            x = g.act(f"pool{bi + 2}_relu", "relu", x)
            x = g.conv(f"pool{bi + 2}_conv", t.conv, x)
            x = g.pool("AveragePooling2D", f"pool{bi + 2}_pool", 2, 2, "valid", x)
    return outs[1:]


def _submodel(name: str, sub, input_name: str, width: int, prefix: str, values: int, sigmoid: bool) -> dict:
    g = Graph(name)
    x = g.input(input_name, (None, None, width))
    for i, c in enumerate(sub.tower):
        x = g.conv(f"{prefix}_{i}", c, x, padding="same", activation="relu", kernel_init=_NORMAL001)
    bias_init = {"class_name": "PriorProbability", "config": {"probability": 0.01}} if sigmoid else _ZEROS
    x = g.conv(prefix, sub.out, x, padding="same", kernel_init=_NORMAL001, bias_init=bias_init)
    x = g.add("Reshape", f"{prefix}_reshape", {"target_shape": [-1, values]}, x)
    if sigmoid:
        x = g.act(f"{prefix}_sigmoid", "sigmoid", x)
    return g.config([x])


def retinanet_config(model) -> dict:
    """The ``config`` of the training ``Model`` (outputs ``regression`` and ``classification``)."""
    g = Graph("retinanet")
    x = g.input("input_1", (None, None, 3))
    bb = model.backbone
    family = type(bb).__name__
    c3, c4, c5 = {"ResNet": _resnet, "VGG": _vgg, "MobileNet": _mobilenet, "DenseNet": _densenet}[family](g, bb, x)
    f = model.fpn
    p5 = g.conv("C5_reduced", f.C5_reduced, c5, padding="same")
    p5_up = g.add("UpsampleLike", "P5_upsampled", {}, [p5, c4])
    p5 = g.conv("P5", f.P5, p5, padding="same")
    p4 = g.conv("C4_reduced", f.C4_reduced, c4, padding="same")
    p4 = g.add("Add", "P4_merged", {}, [p5_up, p4])
    p4_up = g.add("UpsampleLike", "P4_upsampled", {}, [p4, c3])
    p4 = g.conv("P4", f.P4, p4, padding="same")
    p3 = g.conv("C3_reduced", f.C3_reduced, c3, padding="same")
    p3 = g.add("Add", "P3_merged", {}, [p4_up, p3])
    p3 = g.conv("P3", f.P3, p3, padding="same")
    p6 = g.conv("P6", f.P6, c5, padding="same")
    p7 = g.act("C6_relu", "relu", p6)
    p7 = g.conv("P7", f.P7, p7, padding="same")
    feats = [p3, p4, p5, p6, p7]
    width = f.P3.weight.shape[0]
    outs = []
    for sub_name, sub, inp, prefix, values, sig, out_name in (
            ("regression_submodel", model.regression_submodel, "input_2", "pyramid_regression", 4, False, "regression"),
            ("classification_submodel", model.classification_submodel, "input_3", "pyramid_classification",
             model.num_classes, True, "classification")):
        cfg = _submodel(sub_name, sub, inp, width, prefix, values, sig)
        # a Model used as a layer: node 0 is its own construction, the calls on P3..P7 are nodes 1..5
        g.layers.append({"name": sub_name, "class_name": "Model", "config": cfg,
                         "inbound_nodes": [[[t[0], t[1], t[2], {}]] for t in feats]})
        g.index[sub_name] = len(g.layers) - 1
        outs.append(g.add("Concatenate", out_name, {"axis": 1}, [(sub_name, i + 1, 0) for i in range(len(feats))]))
    return g.config(outs)
