"""Synthetic COCO-shaped data: an in-memory generator for benchmarks (every BASELINE.json config
is synthetic) and a writer of tiny on-disk COCO / CSV / VOC / KITTI datasets for tests."""

from __future__ import annotations

import json
import os

import cv2
import numpy as np

from .generator import Generator


class SyntheticGenerator(Generator):
    """Random rectangles on noise; deterministic per index."""

    def __init__(self, num_images=8, image_hw=(512, 512), classes=80, boxes_per_image=6, seed=0, **kwargs):
        self.n, self.hw, self.k, self.nb, self.seed0 = num_images, image_hw, classes, boxes_per_image, seed
        super().__init__(**kwargs)

    def size(self):
        return self.n

    def num_classes(self):
        return self.k

    def name_to_label(self, name):
        return int(name.split("_")[1])

    def label_to_name(self, label):
        return f"class_{label}"

    def image_aspect_ratio(self, image_index):
        return self.hw[1] / self.hw[0]

    def _rng(self, i):
        return np.random.RandomState(self.seed0 * 100003 + i)

    def load_annotations(self, image_index):
        r = self._rng(image_index)
        h, w = self.hw
        out = np.zeros((self.nb, 5))
        for j in range(self.nb):
            bw, bh = r.uniform(0.08, 0.4) * w, r.uniform(0.08, 0.4) * h
            x1, y1 = r.uniform(0, w - bw - 1), r.uniform(0, h - bh - 1)
            out[j] = [x1, y1, x1 + bw, y1 + bh, r.randint(0, self.k)]
        return out

    def load_image(self, image_index):
        r = self._rng(image_index)
        h, w = self.hw
        img = r.randint(90, 150, (h, w, 3)).astype(np.uint8)
        for b in self.load_annotations(image_index):
            color = [int(40 + 25 * ((int(b[4]) + c) % 8)) for c in range(3)]
            cv2.rectangle(img, (int(b[0]), int(b[1])), (int(b[2]), int(b[3])), color, -1)
        return img


def write_coco_dataset(root, set_names=("train2017", "val2017"), num_images=8, image_hw=(128, 160), classes=3, seed=0):
    """Tiny COCO-layout dataset: annotations/instances_<set>.json + images/<set>/*.jpg."""
    gen = SyntheticGenerator(num_images=num_images, image_hw=image_hw, classes=classes, boxes_per_image=3, seed=seed)
    os.makedirs(os.path.join(root, "annotations"), exist_ok=True)
    for s_i, s in enumerate(set_names):
        os.makedirs(os.path.join(root, "images", s), exist_ok=True)
        images, anns = [], []
        for i in range(num_images):
            idx = i + s_i * 1000
            g = SyntheticGenerator(num_images=1, image_hw=image_hw, classes=classes, boxes_per_image=3, seed=seed + idx)
            fn = f"{idx:012d}.jpg"
            cv2.imwrite(os.path.join(root, "images", s, fn), g.load_image(0))
            images.append({"id": idx + 1, "file_name": fn, "height": image_hw[0], "width": image_hw[1]})
            for b in g.load_annotations(0):
                anns.append({"id": len(anns) + 1, "image_id": idx + 1, "category_id": int(b[4]) * 2 + 1,
                             "bbox": [float(b[0]), float(b[1]), float(b[2] - b[0]), float(b[3] - b[1])],
                             "area": float((b[2] - b[0]) * (b[3] - b[1])), "iscrowd": 0})
        cats = [{"id": c * 2 + 1, "name": f"class_{c}", "supercategory": "thing"} for c in range(classes)]
        with open(os.path.join(root, "annotations", f"instances_{s}.json"), "w") as f:
            json.dump({"images": images, "annotations": anns, "categories": cats}, f)
    return root
