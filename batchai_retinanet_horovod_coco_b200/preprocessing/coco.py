"""COCO-format generator (keras-retinanet ``preprocessing/coco.py``; reference ``train.py:197-214``;
SURVEY.md B14).  pycocotools is not available, so the instances JSON is parsed directly:
``<dir>/annotations/instances_<set>.json`` and ``<dir>/images/<set>/<file_name>`` (the code's
layout; the reference README's file names differ -- SURVEY.md Appendix D item 10)."""

from __future__ import annotations

import json
import os

import numpy as np

from ..utils.image import read_image_bgr
from .generator import Generator


class CocoGenerator(Generator):
    def __init__(self, data_dir, set_name, **kwargs):
        self.data_dir, self.set_name = data_dir, set_name
        path = os.path.join(data_dir, "annotations", f"instances_{set_name}.json")
        with open(path) as f:
            self.coco = json.load(f)
        self.images = {im["id"]: im for im in self.coco["images"]}
        self.image_ids = sorted(self.images)
        self.anns_by_image = {}
        for a in self.coco.get("annotations", []):
            self.anns_by_image.setdefault(a["image_id"], []).append(a)
        self.load_classes()
        super().__init__(**kwargs)

    def load_classes(self):
        categories = sorted(self.coco["categories"], key=lambda x: x["id"])
        self.classes, self.coco_labels, self.coco_labels_inverse = {}, {}, {}
        for c in categories:
            self.coco_labels[len(self.classes)] = c["id"]
            self.coco_labels_inverse[c["id"]] = len(self.classes)
            self.classes[c["name"]] = len(self.classes)
        self.labels = {v: k for k, v in self.classes.items()}

    def size(self):
        return len(self.image_ids)

    def num_classes(self):
        return len(self.classes)

    def name_to_label(self, name):
        return self.classes[name]

    def label_to_name(self, label):
        return self.labels[label]

    def coco_label_to_label(self, coco_label):
        return self.coco_labels_inverse[coco_label]

    def coco_label_to_name(self, coco_label):
        return self.label_to_name(self.coco_label_to_label(coco_label))

    def label_to_coco_label(self, label):
        return self.coco_labels[label]

    def image_aspect_ratio(self, image_index):
        im = self.images[self.image_ids[image_index]]
        return float(im["width"]) / float(im["height"])

    def image_path(self, image_index):
        return os.path.join(self.data_dir, "images", self.set_name, self.images[self.image_ids[image_index]]["file_name"])

    def load_image(self, image_index):
        return read_image_bgr(self.image_path(image_index))

    def load_annotations(self, image_index):
        anns = self.anns_by_image.get(self.image_ids[image_index], [])
        out = np.zeros((0, 5))
        for a in anns:
            if a["bbox"][2] < 1 or a["bbox"][3] < 1:  # some annotations have basically no width / height
                continue
            x, y, w, h = a["bbox"]
            out = np.append(out, [[x, y, x + w, y + h, self.coco_label_to_label(a["category_id"])]], axis=0)
        return out
