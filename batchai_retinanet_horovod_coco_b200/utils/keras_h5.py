"""Keras ``.h5`` checkpoint layout: ``load_weights(by_name, skip_mismatch)``, weights-only save,
full-model save (architecture JSON + weights + optimizer config/slots) and full-model load.

Layout per SURVEY.md Appendix B (Keras 2.1/2.2, reference ``train.py:77,113-114,406``):

    /                      attrs keras_version, backend, model_config, training_config
    /model_weights         attrs layer_names, backend, keras_version
    /model_weights/<layer> attrs weight_names; datasets <weight_name> (conv kernels HWIO)
    /optimizer_weights     attrs weight_names; Adam iterations then m_i, v_i per trainable weight

Weights-only files (e.g. ``ResNet-50-model.keras.h5``) carry the ``/model_weights`` content at the
root.  The two shared heads are nested ``Model`` layers (``regression_submodel``,
``classification_submodel``) whose weights sit under one group each.  ``layer_names0..N`` chunked
attributes (newer Keras) are accepted on load.
"""

from __future__ import annotations

import json
import warnings

import numpy as np
import torch

from .h5 import H5File

KERAS_VERSION = b"2.1.6"
BACKEND = b"tensorflow"
SUBMODELS = {"pyramid_regression": "regression_submodel", "pyramid_classification": "classification_submodel"}


def _layer_groups(model):
    """Ordered ``group name -> [(inner layer name, module)]`` as Keras would lay the model out."""
    groups = {}
    for name, mod in model.keras_layers().items():
        g = name
        for prefix, sub in SUBMODELS.items():
            if name == prefix or name.startswith(prefix + "_"):
                g = sub
        groups.setdefault(g, []).append((name, mod))
    return groups


def _is_depthwise(mod) -> bool:
    return getattr(mod, "groups", 1) > 1 and mod.groups == mod.weight.shape[0]


def _to_keras_layout(wname: str, arr):
    if arr.ndim != 4:
        return arr
    return arr.transpose(2, 3, 0, 1) if "depthwise_kernel" in wname else arr.transpose(2, 3, 1, 0)


def _from_keras_layout(wname: str, arr):
    if arr.ndim != 4:
        return arr
    return arr.transpose(2, 3, 0, 1) if "depthwise_kernel" in wname else arr.transpose(3, 2, 0, 1)


def _module_weights(name, mod):
    """[(keras weight name, numpy array in Keras layout, torch tensor)]"""
    out = []
    if hasattr(mod, "moving_variance"):
        for k in ("gamma", "beta", "moving_mean", "moving_variance"):
            t = getattr(mod, k)
            out.append((f"{name}/{k}:0", t.detach().cpu().numpy(), t))
    else:
        w = mod.weight
        if _is_depthwise(mod):   # Keras DepthwiseConv2D: (kh, kw, channels, multiplier)
            out.append((f"{name}/depthwise_kernel:0", w.detach().cpu().permute(2, 3, 0, 1).contiguous().numpy(), w))
        else:
            out.append((f"{name}/kernel:0", w.detach().cpu().permute(2, 3, 1, 0).contiguous().numpy(), w))
        if mod.bias is not None:
            out.append((f"{name}/bias:0", mod.bias.detach().cpu().numpy(), mod.bias))
    return out


def _write_weights(f: H5File, root: str, model):
    groups = _layer_groups(model)
    f.make_group(root)
    f.write_attr(root, "layer_names", [g.encode() for g in groups])
    f.write_attr(root, "backend", BACKEND)
    f.write_attr(root, "keras_version", KERAS_VERSION)
    for g, layers in groups.items():
        path = f"{root.rstrip('/')}/{g}"
        f.make_group(path)
        names = []
        for lname, mod in layers:
            for wname, arr, _ in _module_weights(lname, mod):
                names.append(wname.encode())
                f.write_dataset(f"{path}/{wname}", np.ascontiguousarray(arr, dtype=np.float32))
        f.write_attr(path, "weight_names", names)


def save_weights(model, path: str):
    with H5File(path, "w") as f:
        _write_weights(f, "/", model)


def model_config(model) -> dict:
    """``model.to_json()`` content: a Keras 2.1.6 functional ``Model`` configuration (every layer with its constructor
    arguments and inbound nodes, ``utils/keras_config.py``).  ``b200`` is an extra top-level key Keras ignores; it lets
    ``load_full_model`` rebuild the model without parsing the graph."""
    from .keras_config import retinanet_config
    return {"class_name": "Model", "config": retinanet_config(model), "keras_version": KERAS_VERSION.decode(),
            "backend": BACKEND.decode(), "b200": {"backbone": model.backbone_name, "num_classes": model.num_classes}}


def trainable_weight_list(model):
    """Keras ``model.trainable_weights`` order: layer order, kernel then bias."""
    out = []
    for name, mod in model.keras_layers().items():
        for wname, _, t in _module_weights(name, mod):   # frozen BN vectors are buffers: requires_grad is False
            if t.requires_grad:
                out.append((wname, t))
    return out


def save_model(model, path: str, optimizer_config: dict | None = None, optimizer_slots: dict | None = None,
               iterations: int = 0, epoch: int | None = None):
    """Full-model file as ``ModelCheckpoint`` writes it (``save_weights_only=False``).

    ``optimizer_slots``: {keras weight name: (m, v)} torch/numpy arrays in the **torch** layout
    (OIHW for kernels); they are stored in Keras layout (HWIO) in Keras' order: iterations, all m, all v.
    """
    with H5File(path, "w") as f:
        f.write_attr("/", "keras_version", KERAS_VERSION)
        f.write_attr("/", "backend", BACKEND)
        f.write_attr("/", "model_config", json.dumps(model_config(model), separators=(",", ":")).encode())
        tc = {"optimizer_config": {"class_name": "Adam", "config": optimizer_config or {}},
              "loss": {"regression": "_smooth_l1", "classification": "_focal"}, "metrics": [],
              "sample_weight_mode": None, "loss_weights": None}
        if epoch is not None:
            tc["epoch"] = int(epoch)  # extension: lets --snapshot resume the epoch counter
        f.write_attr("/", "training_config", json.dumps(tc, separators=(",", ":")).encode())
        _write_weights(f, "/model_weights", model)
        if optimizer_slots is not None:
            f.make_group("/optimizer_weights")
            names = [b"Adam/iterations:0"]
            f.write_dataset("/optimizer_weights/Adam/iterations:0", np.asarray(iterations, dtype=np.int64))
            tw = trainable_weight_list(model)
            idx = 0
            for which in (0, 1):
                for wname, t in tw:
                    slot = optimizer_slots.get(wname)
                    arr = np.zeros(tuple(t.shape), np.float32) if slot is None else \
                        np.asarray(slot[which].detach().cpu() if isinstance(slot[which], torch.Tensor) else slot[which],
                                   dtype=np.float32)
                    arr = np.ascontiguousarray(_to_keras_layout(wname, arr))
                    n = "training/Adam/Variable:0" if idx == 0 else f"training/Adam/Variable_{idx}:0"
                    names.append(n.encode())
                    f.write_dataset(f"/optimizer_weights/{n}", arr)
                    idx += 1
            f.write_attr("/optimizer_weights", "weight_names", names)


def _read_names(f: H5File, path: str, attr: str):
    names = f.attr_names(path)
    if attr in names:
        v = f.attr(path, attr)
        return [bytes(x).decode() for x in np.atleast_1d(v)]
    out, i = [], 0
    while f"{attr}{i}" in names:  # chunked attributes of newer Keras
        out += [bytes(x).decode() for x in np.atleast_1d(f.attr(path, f"{attr}{i}"))]
        i += 1
    return out


def load_weights(model, path: str, by_name: bool = True, skip_mismatch: bool = True, verbose: bool = True) -> dict:
    """Name-matched, shape-mismatch-tolerant weight load (reference ``train.py:75-78``).

    Returns ``{"loaded": [...], "skipped": [...], "missing": [...]}`` (weight names).
    """
    layers = dict(model.keras_layers())
    loaded, skipped = [], []
    with H5File(path, "r") as f:
        root = "/model_weights" if f.exists("/model_weights") else "/"
        for g in _read_names(f, root, "layer_names"):
            gpath = f"{root.rstrip('/')}/{g}"
            if not f.exists(gpath):
                continue
            for wname in _read_names(f, gpath, "weight_names"):
                inner, _, leaf = wname.rpartition("/")
                inner = inner.split("/")[-1] if inner else g
                leaf = leaf.split(":")[0]
                mod = layers.get(inner) if inner in layers else layers.get(g)
                if mod is None:
                    continue  # by_name: layers absent from the model are ignored
                arr = f.dataset(f"{gpath}/{wname}")
                t = None
                if leaf in ("kernel", "depthwise_kernel") and hasattr(mod, "weight") and not hasattr(mod, "moving_variance"):
                    t, arr = mod.weight, _from_keras_layout(leaf, arr)
                elif leaf == "bias" and getattr(mod, "bias", None) is not None:
                    t = mod.bias
                elif leaf in ("gamma", "beta", "moving_mean", "moving_variance") and hasattr(mod, leaf):
                    t = getattr(mod, leaf)
                if t is None:
                    continue
                if tuple(arr.shape) != tuple(t.shape):
                    if not skip_mismatch:
                        raise ValueError(f"Layer weight shape {tuple(t.shape)} not compatible with provided weight "
                                         f"shape {tuple(arr.shape)} for {wname}")
                    if verbose:
                        warnings.warn(f"Skipping loading of weights for {wname} due to mismatch in shape "
                                      f"({tuple(t.shape)} vs {tuple(arr.shape)}).")
                    skipped.append(wname)
                    continue
                with torch.no_grad():
                    t.copy_(torch.from_numpy(np.ascontiguousarray(arr)).to(t.dtype))
                loaded.append(wname)
    have = set(loaded) | set(skipped)
    missing = [w for n, m in layers.items() for w, _, _ in _module_weights(n, m) if w not in have]
    return {"loaded": loaded, "skipped": skipped, "missing": missing}


def load_full_model(path: str, backbone_name: str = "resnet50", num_classes: int | None = None):
    """``models.load_model`` (reference ``train.py:406``): architecture + weights + optimizer state.

    Returns ``(model, training_state)`` with ``training_state = {"optimizer_config", "iterations",
    "slots": {weight name: (m, v)}, "epoch"}`` (slots in torch layout).
    """
    from ..models.retinanet import RetinaNet
    state = {"optimizer_config": {}, "iterations": 0, "slots": {}, "epoch": None}
    with H5File(path, "r") as f:
        names = f.attr_names("/")
        cfg = json.loads(bytes(f.attr("/", "model_config")).decode()) if "model_config" in names else {}
        c = cfg.get("b200") or cfg.get("config", {})          # "config" held these keys in files written before round 2
        backbone_name = c.get("backbone", backbone_name)
        if num_classes is None:
            num_classes = c.get("num_classes")
        if num_classes is None:
            k = f.dataset("/model_weights/classification_submodel/pyramid_classification/bias:0")
            num_classes = k.shape[0] // 9
        model = RetinaNet(int(num_classes), backbone_name)
        if "training_config" in names:
            tc = json.loads(bytes(f.attr("/", "training_config")).decode())
            state["optimizer_config"] = tc.get("optimizer_config", {}).get("config", {})
            state["epoch"] = tc.get("epoch")
        if f.exists("/optimizer_weights"):
            wn = _read_names(f, "/optimizer_weights", "weight_names")
            arrs = [f.dataset(f"/optimizer_weights/{n}") for n in wn]
            if arrs:
                state["iterations"] = int(np.asarray(arrs[0]).reshape(-1)[0])
                tw = trainable_weight_list(model)
                rest = arrs[1:]
                if len(rest) == 2 * len(tw):
                    for i, (wname, t) in enumerate(tw):
                        m, v = rest[i], rest[len(tw) + i]
                        m, v = _from_keras_layout(wname, m), _from_keras_layout(wname, v)
                        state["slots"][wname] = (torch.from_numpy(np.ascontiguousarray(m)),
                                                 torch.from_numpy(np.ascontiguousarray(v)))
    load_weights(model, path, by_name=True, skip_mismatch=False, verbose=False)
    return model, state
