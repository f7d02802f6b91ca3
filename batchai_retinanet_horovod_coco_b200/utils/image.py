"""Image IO and geometry (keras-retinanet ``utils/image.py``):
BGR read, Caffe-mode preprocessing (subtract 103.939/116.779/123.68, no scaling), resize so the
short side is ``min_side`` capped so the long side stays <= ``max_side``, affine warps."""

from __future__ import annotations

import cv2
import numpy as np

from .transform import change_transform_origin

CAFFE_MEAN_BGR = np.array([103.939, 116.779, 123.68], dtype=np.float32)


def read_image_bgr(path: str) -> np.ndarray:
    img = cv2.imread(path, cv2.IMREAD_COLOR)
    if img is None:
        raise FileNotFoundError(f"cannot read image {path}")
    return img


def preprocess_image(x: np.ndarray, mode: str = "caffe") -> np.ndarray:
    """``caffe``: subtract the BGR ImageNet mean (ResNet, VGG); ``tf``: scale to [-1, 1] (MobileNet, DenseNet)."""
    x = x.astype(np.float32)
    if mode == "tf":
        return x / 127.5 - 1.0
    if mode != "caffe":
        raise ValueError(f"unknown preprocessing mode '{mode}'")
    return x - CAFFE_MEAN_BGR


def compute_resize_scale(image_shape, min_side=800, max_side=1333) -> float:
    rows, cols = image_shape[0], image_shape[1]
    smallest, largest = min(rows, cols), max(rows, cols)
    scale = min_side / smallest
    if largest * scale > max_side:
        scale = max_side / largest
    return scale


def resize_image(img: np.ndarray, min_side=800, max_side=1333):
    scale = compute_resize_scale(img.shape, min_side, max_side)
    return cv2.resize(img, None, fx=scale, fy=scale), scale


class TransformParameters:
    def __init__(self, fill_mode="nearest", interpolation="linear", cval=0, relative_translation=True):
        self.fill_mode, self.cval, self.interpolation = fill_mode, cval, interpolation
        self.relative_translation = relative_translation

    def cv_border_mode(self):
        return {"constant": cv2.BORDER_CONSTANT, "nearest": cv2.BORDER_REPLICATE, "reflect": cv2.BORDER_REFLECT_101,
                "wrap": cv2.BORDER_WRAP}[self.fill_mode]

    def cv_interpolation(self):
        return {"nearest": cv2.INTER_NEAREST, "linear": cv2.INTER_LINEAR, "cubic": cv2.INTER_CUBIC,
                "area": cv2.INTER_AREA, "lanczos4": cv2.INTER_LANCZOS4}[self.interpolation]


def adjust_transform_for_image(transform, image, relative_translation):
    """Move the transform origin to the image centre; scale translation by the image size."""
    height, width = image.shape[:2]
    result = transform.copy()
    if relative_translation:
        result[0:2, 2] *= [width, height]
    return change_transform_origin(result, (0.5 * width, 0.5 * height))


def apply_transform(matrix, image, params: TransformParameters):
    return cv2.warpAffine(image, matrix[:2, :], dsize=(image.shape[1], image.shape[0]), flags=params.cv_interpolation(),
                          borderMode=params.cv_border_mode(), borderValue=params.cval)
