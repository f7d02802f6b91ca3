"""Open Images generator (keras-retinanet ``preprocessing/open_images.py``; reference
``train.py:252-276``): ``<main_dir>/<version>/class-descriptions-boxable.csv``,
``<subset>-annotations-bbox.csv`` (normalised XMin/XMax/YMin/YMax), images under
``<main_dir>/images/<subset>/<ImageID>.jpg``; a JSON annotation cache in ``annotation_cache_dir``;
optional ``labels_filter`` (keep only these class names; ``fixed_labels`` keeps their given order)."""

from __future__ import annotations

import csv
import json
import os

import numpy as np
from PIL import Image

from ..utils.image import read_image_bgr
from .generator import Generator


def get_labels(metadata_dir, version="v4"):
    path = os.path.join(metadata_dir, "class-descriptions-boxable.csv" if version == "v4" else "classes-bbox.txt")
    id_to_labels, cls_index = {}, {}
    with open(path, newline="") as f:
        for i, row in enumerate(csv.reader(f)):
            if not row:
                continue
            label = row[0]
            id_to_labels[i] = row[1] if len(row) > 1 else label
            cls_index[label] = i
    return id_to_labels, cls_index


def generate_images_annotations_json(main_dir, metadata_dir, subset, cls_index, version="v4"):
    path = os.path.join(metadata_dir, subset, f"{subset}-annotations-bbox.csv")
    if not os.path.exists(path):
        path = os.path.join(metadata_dir, f"{subset}-annotations-bbox.csv")
    id_annotations = {}
    with open(path, newline="") as f:
        reader = csv.DictReader(f)
        for row in reader:
            frame = row["ImageID"]
            if row["LabelName"] not in cls_index:
                continue
            img_path = os.path.join(main_dir, "images", subset, frame + ".jpg")
            entry = id_annotations.get(frame)
            if entry is None:
                if not os.path.exists(img_path):
                    continue
                with Image.open(img_path) as im:
                    w, h = im.size
                entry = id_annotations[frame] = {"w": w, "h": h, "boxes": []}
            x1, x2, y1, y2 = float(row["XMin"]), float(row["XMax"]), float(row["YMin"]), float(row["YMax"])
            x1i, x2i = int(round(x1 * entry["w"])), int(round(x2 * entry["w"]))
            y1i, y2i = int(round(y1 * entry["h"])), int(round(y2 * entry["h"]))
            if x2i <= x1i or y2i <= y1i:
                continue  # degenerate after rounding
            entry["boxes"].append({"cls_id": cls_index[row["LabelName"]], "x1": x1, "x2": x2, "y1": y1, "y2": y2})
    return id_annotations


class OpenImagesGenerator(Generator):
    def __init__(self, main_dir, subset, version="v4", labels_filter=None, annotation_cache_dir=".", fixed_labels=False,
                 **kwargs):
        metadata = "challenge2018" if version == "challenge2018" else "2018_04" if version == "v4" else "2017_11"
        self.base_dir = os.path.join(main_dir, "images", subset)
        metadata_dir = os.path.join(main_dir, metadata)
        if not os.path.isdir(metadata_dir):
            metadata_dir = os.path.join(main_dir, version)
        cache = os.path.join(annotation_cache_dir, subset + ".json")
        self.id_to_labels, cls_index = get_labels(metadata_dir, version)
        if os.path.exists(cache):
            with open(cache) as f:
                self.annotations = json.load(f)
        else:
            self.annotations = generate_images_annotations_json(main_dir, metadata_dir, subset, cls_index, version)
            with open(cache, "w") as f:
                json.dump(self.annotations, f)
        if labels_filter is not None:
            self.id_to_labels, self.annotations = self._filter(labels_filter, fixed_labels)
        self.id_to_image_id = dict(enumerate(self.annotations))
        super().__init__(**kwargs)

    def _filter(self, labels_filter, fixed_labels):
        labels_to_id = {l: i for i, l in self.id_to_labels.items()}
        sub_labels_to_id = {}
        if fixed_labels:
            sub_labels_to_id = {l: i for i, l in enumerate(labels_filter)}
        else:
            for l in sorted(labels_filter, key=lambda x: labels_to_id[x]):
                sub_labels_to_id[l] = len(sub_labels_to_id)
        id_map = {labels_to_id[l]: i for l, i in sub_labels_to_id.items()}
        filtered = {}
        for k, ann in self.annotations.items():
            boxes = [dict(b, cls_id=id_map[b["cls_id"]]) for b in ann["boxes"] if b["cls_id"] in id_map]
            if boxes:
                filtered[k] = {"w": ann["w"], "h": ann["h"], "boxes": boxes}
        return {i: l for l, i in sub_labels_to_id.items()}, filtered

    def size(self):
        return len(self.annotations)

    def num_classes(self):
        return len(self.id_to_labels)

    def name_to_label(self, name):
        raise NotImplementedError()

    def label_to_name(self, label):
        return self.id_to_labels[label]

    def image_aspect_ratio(self, image_index):
        a = self.annotations[self.id_to_image_id[image_index]]
        return float(a["w"]) / float(a["h"])

    def image_path(self, image_index):
        return os.path.join(self.base_dir, self.id_to_image_id[image_index] + ".jpg")

    def load_image(self, image_index):
        return read_image_bgr(self.image_path(image_index))

    def load_annotations(self, image_index):
        a = self.annotations[self.id_to_image_id[image_index]]
        boxes = np.zeros((len(a["boxes"]), 5))
        for idx, b in enumerate(a["boxes"]):
            boxes[idx] = [b["x1"] * a["w"], b["y1"] * a["h"], b["x2"] * a["w"], b["y2"] * a["h"], b["cls_id"]]
        return boxes
