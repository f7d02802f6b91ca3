"""Writes tests/fixtures/keras_resnet50_stem_libhdf5.h5: the first layers of a keras-retinanet ResNet-50 weight file
(`model.save_weights`, Keras 2.1.6 / TensorFlow backend) in libhdf5 1.8's byte layout -- see libhdf5_layout.py.
Values are a fixed pseudo-random stream so the test can recompute them.  Run from the repo root to regenerate."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from libhdf5_layout import LibHdf5File  # noqa: E402

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "keras_resnet50_stem_libhdf5.h5")

LAYERS = [  # (layer, [(weight name, shape)])  -- in Keras' layer order; layers without weights are listed too
    ("input_1", []),
    ("padding_conv1", []),
    ("conv1", [("conv1/kernel:0", (7, 7, 3, 64))]),
    ("bn_conv1", [("bn_conv1/gamma:0", (64,)), ("bn_conv1/beta:0", (64,)), ("bn_conv1/moving_mean:0", (64,)),
                  ("bn_conv1/moving_variance:0", (64,))]),
    ("conv1_relu", []),
    ("pool1", []),
    ("res2a_branch2a", [("res2a_branch2a/kernel:0", (1, 1, 64, 64))]),
    ("bn2a_branch2a", [("bn2a_branch2a/gamma:0", (64,)), ("bn2a_branch2a/beta:0", (64,)),
                       ("bn2a_branch2a/moving_mean:0", (64,)), ("bn2a_branch2a/moving_variance:0", (64,))]),
    # a layer this package's models do not have (by_name loading must ignore it) and one with a foreign shape
    ("fc1000", [("fc1000/kernel:0", (8, 10)), ("fc1000/bias:0", (10,))]),
    ("res2a_branch2b", [("res2a_branch2b/kernel:0", (3, 3, 64, 32))]),
]


def values(name: str, shape):
    seed = sum(ord(c) * (i + 1) for i, c in enumerate(name)) % (2 ** 31)
    v = np.random.RandomState(seed).standard_normal(int(np.prod(shape))).astype(np.float32).reshape(shape)
    return np.abs(v) + 0.5 if name.endswith("moving_variance:0") else v


def build() -> bytes:
    f = LibHdf5File()
    f.attr("/", "layer_names", np.array([n.encode() for n, _ in LAYERS]))
    f.attr("/", "backend", b"tensorflow")            # Python bytes -> variable-length strings in the global heap
    f.attr("/", "keras_version", b"2.1.6")
    for layer, weights in LAYERS:
        f.group("/" + layer)
        if weights:
            f.attr("/" + layer, "weight_names", np.array([w.encode() for w, _ in weights]))
        else:
            f.attr("/" + layer, "weight_names", np.zeros((0,), dtype="S1"))
        for w, shape in weights:
            f.dataset(f"/{layer}/{w}", values(w, shape))
    return f.tobytes()


if __name__ == "__main__":
    with open(PATH, "wb") as fh:
        fh.write(build())
    print(PATH, os.path.getsize(PATH), "bytes")
