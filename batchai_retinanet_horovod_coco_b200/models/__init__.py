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


class UnsupportedBackbone(Backbone):
    def validate(self):
        raise NotImplementedError(
            f"Backbone '{self.backbone}' is recognised by the reference CLI (train.py:323-324 warns it is "
            f"experimental) but only resnet50/101/152 are implemented here.")


def backbone(name: str) -> Backbone:
    if "resnet" in name:
        return Backbone(name)
    if any(k in name for k in ("mobilenet", "vgg", "densenet")):
        return UnsupportedBackbone(name)
    raise NotImplementedError(f"Backbone class for '{name}' not implemented.")


def load_model(filepath: str, backbone_name: str = "resnet50", num_classes: int | None = None):
    """Full-model load (architecture + weights + optimizer state) from a Keras-layout ``.h5``.

    Reference ``train.py:406``.  Returns ``(model, training_state)`` where ``training_state``
    holds the optimizer config/slots and epoch counter if the file has them.
    """
    from ..utils.keras_h5 import load_full_model
    return load_full_model(filepath, backbone_name=backbone_name, num_classes=num_classes)
