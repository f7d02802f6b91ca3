"""Dataset generator base class (keras-retinanet ``preprocessing/generator.py``; built by
``create_generators`` at reference ``train.py:177-297``).

Same contract as upstream -- infinite, thread-safe iterator; images grouped by aspect ratio and
group order reshuffled every pass; invalid boxes dropped with a warning; optional random affine;
resize to min/max side; zero-padding to the batch maximum (top-left aligned) -- with two output
forms:

* ``next()``        Keras form: ``(images float32 [B,H,W,3] mean-subtracted, [regression (B,A,5),
                    labels (B,A,K+1)])`` -- what the reference feeds ``fit_generator``;
* ``next_compact()`` engine form: ``(images uint8 [B,H,W,3], gt float32 [B,G,5], gt_count int32 [B],
                    img_hw int32 [B,2])`` -- anchor targets are assigned on the GPU instead of
                    shipping 65 MB/image of one-hot labels over PCIe.

Differences from upstream, both opt-in: ``seed`` (upstream uses the global unseeded ``random``)
and ``shard=(rank, size)`` to give every rank a disjoint slice.
"""

from __future__ import annotations

import random
import threading
import warnings

import numpy as np
import torch

from ..ops.anchors import anchor_targets_bbox, anchors_for_shape
from ..utils.image import (TransformParameters, adjust_transform_for_image, apply_transform, compute_resize_scale,
                           preprocess_image,
                           resize_image)
from ..utils.transform import transform_aabb

PAD_VALUE_BGR = (104, 117, 124)  # uint8 padding that becomes ~0 after Caffe mean subtraction


class Generator:
    def __init__(self, transform_generator=None, batch_size=1, group_method="ratio", shuffle_groups=True,
                 image_min_side=800, image_max_side=1333, transform_parameters=None, compute_anchor_targets=None,
                 seed=None, shard=None):
        self.transform_generator = transform_generator
        self.batch_size = int(batch_size)
        self.group_method = group_method
        self.shuffle_groups = shuffle_groups
        self.image_min_side = image_min_side
        self.image_max_side = image_max_side
        self.transform_parameters = transform_parameters or TransformParameters()
        self.compute_anchor_targets = compute_anchor_targets or anchor_targets_bbox
        self.shapes_callback = None
        self.rng = random.Random(seed) if seed is not None else random
        self.shard = shard
        self.group_index = 0
        self.lock = threading.Lock()
        self.group_images()

    # ---- dataset interface (subclasses) --------------------------------------------------
    def size(self):
        raise NotImplementedError

    def num_classes(self):
        raise NotImplementedError

    def name_to_label(self, name):
        raise NotImplementedError

    def label_to_name(self, label):
        raise NotImplementedError

    def image_aspect_ratio(self, image_index):
        raise NotImplementedError

    def load_image(self, image_index):
        raise NotImplementedError

    def load_annotations(self, image_index):
        """(n,5) float array x1,y1,x2,y2,label."""
        raise NotImplementedError

    # ---- pipeline -------------------------------------------------------------------------
    def load_annotations_group(self, group):
        return [np.asarray(self.load_annotations(i), dtype=np.float64).reshape(-1, 5) for i in group]

    def filter_annotations(self, image_group, annotations_group, group):
        for index, (image, annotations) in enumerate(zip(image_group, annotations_group)):
            a = annotations
            invalid = np.where((a[:, 2] <= a[:, 0]) | (a[:, 3] <= a[:, 1]) | (a[:, 0] < 0) | (a[:, 1] < 0) |
                               (a[:, 2] > image.shape[1]) | (a[:, 3] > image.shape[0]))[0]
            if len(invalid):
                warnings.warn(f"Image with id {group[index]} (shape {image.shape}) contains the following invalid "
                              f"boxes: {a[invalid, :].tolist()}.")
                annotations_group[index] = np.delete(a, invalid, axis=0)
        return image_group, annotations_group

    def load_image_group(self, group):
        return [self.load_image(i) for i in group]

    def random_transform_group_entry(self, image, annotations):
        if self.transform_generator:
            with self.lock:      # Python generators are not re-entrant; prefetch workers share this one
                matrix = next(self.transform_generator)
            transform = adjust_transform_for_image(matrix, image,
                                                   self.transform_parameters.relative_translation)
            image = apply_transform(transform, image, self.transform_parameters)
            annotations = annotations.copy()
            for index in range(annotations.shape[0]):
                annotations[index, :4] = transform_aabb(transform, annotations[index, :4])
        return image, annotations

    def resize_image(self, image):
        return resize_image(image, min_side=self.image_min_side, max_side=self.image_max_side)

    def preprocess_group_entry(self, image, annotations, keras_form):
        image, annotations = self.random_transform_group_entry(image, annotations)
        if keras_form:
            image = preprocess_image(image)
        image, scale = self.resize_image(image)
        annotations = annotations.copy()
        annotations[:, :4] *= scale
        return image, annotations

    def group_images(self):
        order = list(range(self.size()))
        if self.shard is not None:
            order = order[self.shard[0]::self.shard[1]]
        if self.group_method == "random":
            self.rng.shuffle(order)
        elif self.group_method == "ratio":
            order.sort(key=lambda x: self.image_aspect_ratio(x))
        b = self.batch_size
        # last group wraps around so that every group is full (upstream behaviour)
        self.groups = [[order[x % len(order)] for x in range(i, i + b)] for i in range(0, len(order), b)]

    def _next_group(self):
        with self.lock:
            if self.group_index == 0 and self.shuffle_groups:
                self.rng.shuffle(self.groups)
            group = self.groups[self.group_index]
            self.group_index = (self.group_index + 1) % len(self.groups)
        return group

    def _load_group(self, group, keras_form):
        image_group = self.load_image_group(group)
        annotations_group = self.load_annotations_group(group)
        image_group, annotations_group = self.filter_annotations(image_group, annotations_group, group)
        for i in range(len(image_group)):
            image_group[i], annotations_group[i] = self.preprocess_group_entry(image_group[i], annotations_group[i],
                                                                               keras_form)
        return image_group, annotations_group

    def compute_inputs(self, image_group, dtype=np.float32, fill=0):
        max_shape = tuple(max(image.shape[x] for image in image_group) for x in range(3))
        batch = np.empty((len(image_group),) + max_shape, dtype=dtype)
        batch[...] = np.asarray(fill, dtype=dtype)
        for i, image in enumerate(image_group):
            batch[i, :image.shape[0], :image.shape[1], :image.shape[2]] = image
        return batch

    def compute_targets(self, image_group, annotations_group):
        max_shape = tuple(max(image.shape[x] for image in image_group) for x in range(2))
        anchors = anchors_for_shape(max_shape, shapes_callback=self.shapes_callback)
        reg, lab = self.compute_anchor_targets(anchors, annotations_group, self.num_classes(),
                                               [im.shape[:2] for im in image_group])
        return [reg.numpy(), lab.numpy()]

    # ---- iteration ------------------------------------------------------------------------
    def __iter__(self):
        return self

    def __next__(self):
        return self.next()

    def next(self):
        image_group, annotations_group = self._load_group(self._next_group(), keras_form=True)
        return self.compute_inputs(image_group), self.compute_targets(image_group, annotations_group)

    def next_raw(self):
        """The batch BEFORE any pixel work: decoded images, augmentation matrices, resize scales, transformed boxes.
        ``preprocessing.gpu_batch.GpuBatchAssembler`` turns it into the engine's compact batch with the warp and the
        resize running on the GPU (same OpenCV arithmetic), so worker threads only decode."""
        group = self._next_group()
        image_group = self.load_image_group(group)
        annotations_group = self.load_annotations_group(group)
        image_group, annotations_group = self.filter_annotations(image_group, annotations_group, group)
        matrices, out_hw, scales = [], [], []
        for i, (image, ann) in enumerate(zip(image_group, annotations_group)):
            ann = ann.copy()
            m = None
            if self.transform_generator:
                with self.lock:
                    matrix = next(self.transform_generator)
                m = adjust_transform_for_image(matrix, image, self.transform_parameters.relative_translation)
                for k in range(ann.shape[0]):
                    ann[k, :4] = transform_aabb(m, ann[k, :4])
            scale = compute_resize_scale(image.shape, self.image_min_side, self.image_max_side)
            # cv2.resize(img, None, fx=scale, fy=scale): dsize = round(size * scale); source step = 1 / scale
            h, w = int(round(image.shape[0] * scale)), int(round(image.shape[1] * scale))
            ann[:, :4] *= scale
            annotations_group[i] = ann
            matrices.append(m)
            out_hw.append((h, w))
            scales.append((1.0 / scale, 1.0 / scale))
        b = len(image_group)
        g = max(1, max(a.shape[0] for a in annotations_group))
        gt = torch.zeros(b, g, 5)
        cnt = torch.zeros(b, dtype=torch.int32)
        for i, a in enumerate(annotations_group):
            gt[i, :a.shape[0]] = torch.from_numpy(a.astype(np.float32))
            cnt[i] = a.shape[0]
        tp = self.transform_parameters
        return {"images": image_group, "matrices": matrices, "out_hw": out_hw, "scales": scales, "gt": gt, "count": cnt,
                "border_replicate": tp.fill_mode == "nearest", "cval": int(tp.cval)}

    def next_compact(self, pin: bool = False):
        image_group, annotations_group = self._load_group(self._next_group(), keras_form=False)
        images = torch.from_numpy(self.compute_inputs(image_group, dtype=np.uint8, fill=PAD_VALUE_BGR))
        b = len(image_group)
        g = max(1, max(a.shape[0] for a in annotations_group))
        gt = torch.zeros(b, g, 5)
        cnt = torch.zeros(b, dtype=torch.int32)
        for i, a in enumerate(annotations_group):
            gt[i, :a.shape[0]] = torch.from_numpy(a.astype(np.float32))
            cnt[i] = a.shape[0]
        hw = torch.tensor([[im.shape[0], im.shape[1]] for im in image_group], dtype=torch.int32)
        out = (images, gt, cnt, hw)
        if pin and torch.cuda.is_available():
            out = tuple(t.pin_memory() for t in out)
        return out
