"""KITTI generator (keras-retinanet ``preprocessing/kitti.py``; reference ``train.py:277-293``):
``<dir>/<subset>/labels/*.txt`` (KITTI object label rows) and ``<dir>/<subset>/images/*.png``."""

from __future__ import annotations

import csv
import os

import numpy as np
from PIL import Image

from ..utils.image import read_image_bgr
from .generator import Generator

kitti_classes = {"Car": 0, "Van": 1, "Truck": 2, "Pedestrian": 3, "Person_sitting": 4, "Cyclist": 5, "Tram": 6,
                 "Misc": 7, "DontCare": 7}


class KittiGenerator(Generator):
    def __init__(self, base_dir, subset="train", **kwargs):
        self.base_dir = base_dir
        label_dir = os.path.join(base_dir, subset, "labels")
        image_dir = os.path.join(base_dir, subset, "images")
        self.classes = kitti_classes
        self.labels = {}
        for name, label in self.classes.items():
            self.labels.setdefault(label, name)
        self.image_data, self.images = {}, []
        fields = ["type", "truncated", "occluded", "alpha", "left", "top", "right", "bottom", "dh", "dw", "dl", "lx",
                  "ly", "lz", "ry"]
        for i, fn in enumerate(sorted(os.listdir(label_dir))):
            self.images.append(os.path.join(image_dir, fn.replace(".txt", ".png")))
            boxes = []
            with open(os.path.join(label_dir, fn)) as f:
                for line, row in enumerate(csv.DictReader(f, delimiter=" ", fieldnames=fields)):
                    cls_id = self.classes.get(row["type"])
                    if cls_id is None:
                        raise ValueError(f"{fn} line {line}: unknown KITTI class {row['type']}")
                    boxes.append({"cls_id": cls_id, "x1": float(row["left"]), "x2": float(row["right"]),
                                  "y1": float(row["top"]), "y2": float(row["bottom"])})
            self.image_data[i] = boxes
        super().__init__(**kwargs)

    def size(self):
        return len(self.images)

    def num_classes(self):
        return max(self.classes.values()) + 1

    def name_to_label(self, name):
        return self.classes[name]

    def label_to_name(self, label):
        return self.labels[label]

    def image_aspect_ratio(self, image_index):
        with Image.open(self.images[image_index]) as im:
            return float(im.width) / float(im.height)

    def load_image(self, image_index):
        return read_image_bgr(self.images[image_index])

    def load_annotations(self, image_index):
        ann = self.image_data[image_index]
        boxes = np.zeros((len(ann), 5))
        for idx, a in enumerate(ann):
            boxes[idx] = [a["x1"], a["y1"], a["x2"], a["y2"], a["cls_id"]]
        return boxes
