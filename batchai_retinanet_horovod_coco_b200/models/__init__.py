"""Model factory: ``backbone(name)`` -> object with ``.retinanet(...)``, ``.download_imagenet()``,
``.validate()``; ``load_model(path, backbone_name)``.

Mirrors the keras-retinanet ``models`` surface the reference binds at ``train.py:43,390,406,413``
(SURVEY.md B1, B4, B21).
"""

from __future__ import annotations

import os
import warnings

from .resnet import BLOCKS, ResNet, FrozenBatchNorm2d
from .retinanet import RetinaNet, RetinaNetBBox, retinanet_bbox

IMAGENET_FILES = {
    "resnet50": "ResNet-50-model.keras.h5",
    "resnet101": "ResNet-101-model.keras.h5",
    "resnet152": "ResNet-152-model.keras.h5",
}


class Backbone:
    """Stores backbone information (reference ``train.py:390``)."""

    def __init__(self, name: str):
        self.backbone = name
        self.validate()

    def retinanet(self, num_classes: int, modifier=None, **kw) -> RetinaNet:
        model = RetinaNet(num_classes=num_classes, backbone=self.backbone, **kw)
        if modifier is not None:
            model = modifier(model) or model
        return model

    def download_imagenet(self) -> str | None:
        """Resolve the ImageNet weight file **locally** (there is no network on a B200 box).

        Search order: ``$RETINANET_IMAGENET_WEIGHTS`` (file or directory), ``~/.keras/models``.
        Returns ``None`` with a warning when absent, which degrades to ``--no-weights``.
        """
        fname = IMAGENET_FILES[self.backbone]
        cands = []
        env = os.environ.get("RETINANET_IMAGENET_WEIGHTS")
        if env:
            cands += [env, os.path.join(env, fname)]
        cands.append(os.path.join(os.path.expanduser("~"), ".keras", "models", fname))
        for c in cands:
            if os.path.isfile(c):
                return c
        warnings.warn(f"ImageNet weights '{fname}' not found locally and there is no network; "
                      f"continuing with random initialisation (same as --no-weights). "
                      f"Set RETINANET_IMAGENET_WEIGHTS to a file or directory to use them.")
        return None

    def validate(self):
        if self.backbone not in BLOCKS:
            raise ValueError(f"Backbone ('{self.backbone}') not in allowed backbones ({sorted(BLOCKS)}).")

    def preprocess_image(self, x):
        """caffe mode: BGR mean subtraction (keras-retinanet ``preprocess_image``)."""
        from ..utils.image import preprocess_image
        return preprocess_image(x, mode="caffe")


class VGGBackbone(Backbone):
    def validate(self):
        from .vgg import VGG_CFG
        if self.backbone not in VGG_CFG:
            raise ValueError(f"Backbone ('{self.backbone}') not in allowed backbones ({sorted(VGG_CFG)}).")

    def download_imagenet(self):
        return _local_weights(f"{self.backbone}_weights_tf_dim_ordering_tf_kernels_notop.h5")


class MobileNetBackbone(Backbone):
    def validate(self):
        from .mobilenet import ALLOWED
        if self.backbone not in ALLOWED:
            raise ValueError(f"Backbone ('{self.backbone}') not in allowed backbones ({ALLOWED}).")

    def download_imagenet(self):
        res, alpha = self.backbone[len("mobilenet"):].split("_")
        a = {"1.0": "1_0", "0.75": "7_5", "0.5": "5_0", "0.25": "2_5"}[alpha]
        return _local_weights(f"mobilenet_{a}_{res}_tf_no_top.h5")

    def preprocess_image(self, x):
        from ..utils.image import preprocess_image
        return preprocess_image(x, mode="tf")


class DenseNetBackbone(Backbone):
    def validate(self):
        from .densenet import DENSENET_BLOCKS
        if self.backbone not in DENSENET_BLOCKS:
            raise ValueError(f"Backbone ('{self.backbone}') not in allowed backbones ({sorted(DENSENET_BLOCKS)}).")

    def download_imagenet(self):
        return _local_weights(f"{self.backbone}_weights_tf_dim_ordering_tf_kernels_notop.h5")

    def preprocess_image(self, x):
        from ..utils.image import preprocess_image
        return preprocess_image(x, mode="tf")


def _local_weights(fname: str):
    """Keras ``get_file`` without the network: ``$RETINANET_IMAGENET_WEIGHTS`` or ``~/.keras/models``."""
    cands = []
    env = os.environ.get("RETINANET_IMAGENET_WEIGHTS")
    if env:
        cands += [env, os.path.join(env, fname)]
    cands.append(os.path.join(os.path.expanduser("~"), ".keras", "models", fname))
    for c in cands:
        if os.path.isfile(c):
            return c
    warnings.warn(f"ImageNet weights '{fname}' not found locally and there is no network; continuing with random "
                  f"initialisation (same as --no-weights). Set RETINANET_IMAGENET_WEIGHTS to use them.")
    return None


def backbone(name: str) -> Backbone:
    """Dispatch by substring exactly like keras-retinanet (reference ``train.py:43,390``)."""
    if "resnet" in name:
        return Backbone(name)
    if "mobilenet" in name:
        return MobileNetBackbone(name)
    if "vgg" in name:
        return VGGBackbone(name)
    if "densenet" in name:
        return DenseNetBackbone(name)
    raise NotImplementedError(f"Backbone class for '{name}' not implemented.")


def load_model(filepath: str, backbone_name: str = "resnet50", num_classes: int | None = None):
    """Full-model load (architecture + weights + optimizer state) from a Keras-layout ``.h5``.

    Reference ``train.py:406``.  Returns ``(model, training_state)`` where ``training_state``
    holds the optimizer config/slots and epoch counter if the file has them.
    """
    from ..utils.keras_h5 import load_full_model
    return load_full_model(filepath, backbone_name=backbone_name, num_classes=num_classes)
