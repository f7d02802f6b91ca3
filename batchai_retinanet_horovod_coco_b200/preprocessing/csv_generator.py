"""CSV generator (keras-retinanet ``preprocessing/csv_generator.py``; reference ``train.py:232-251``).
Annotations: ``path,x1,y1,x2,y2,class_name`` (empty box fields = image without objects);
classes: ``class_name,id``."""

from __future__ import annotations

import csv
import os
from collections import OrderedDict

import numpy as np
from PIL import Image

from ..utils.image import read_image_bgr
from .generator import Generator


def _parse(value, function, fmt):
    try:
        return function(value)
    except ValueError as e:
        raise ValueError(fmt.format(e)) from None


def read_classes(reader):
    result = OrderedDict()
    for line, row in enumerate(reader, 1):
        try:
            class_name, class_id = row
        except ValueError:
            raise ValueError(f"line {line}: format should be 'class_name,class_id'") from None
        class_id = _parse(class_id, int, f"line {line}: malformed class ID: {{}}")
        if class_name in result:
            raise ValueError(f"line {line}: duplicate class name: '{class_name}'")
        result[class_name] = class_id
    return result


def read_annotations(reader, classes):
    result = OrderedDict()
    for line, row in enumerate(reader, 1):
        try:
            img_file, x1, y1, x2, y2, class_name = row[:6]
        except ValueError:
            raise ValueError(f"line {line}: format should be 'img_file,x1,y1,x2,y2,class_name' or 'img_file,,,,,'") from None
        result.setdefault(img_file, [])
        if (x1, y1, x2, y2, class_name) == ("", "", "", "", ""):
            continue
        x1 = _parse(x1, int, f"line {line}: malformed x1: {{}}")
        y1 = _parse(y1, int, f"line {line}: malformed y1: {{}}")
        x2 = _parse(x2, int, f"line {line}: malformed x2: {{}}")
        y2 = _parse(y2, int, f"line {line}: malformed y2: {{}}")
        if x2 <= x1:
            raise ValueError(f"line {line}: x2 ({x2}) must be higher than x1 ({x1})")
        if y2 <= y1:
            raise ValueError(f"line {line}: y2 ({y2}) must be higher than y1 ({y1})")
        if class_name not in classes:
            raise ValueError(f"line {line}: unknown class name: '{class_name}' (classes: {list(classes)})")
        result[img_file].append({"x1": x1, "x2": x2, "y1": y1, "y2": y2, "class": class_name})
    return result


class CSVGenerator(Generator):
    def __init__(self, csv_data_file, csv_class_file, base_dir=None, **kwargs):
        self.base_dir = base_dir if base_dir is not None else os.path.dirname(csv_data_file)
        with open(csv_class_file, newline="") as f:
            self.classes = read_classes(csv.reader(f, delimiter=","))
        self.labels = {v: k for k, v in self.classes.items()}
        with open(csv_data_file, newline="") as f:
            self.image_data = read_annotations(csv.reader(f, delimiter=","), self.classes)
        self.image_names = list(self.image_data.keys())
        super().__init__(**kwargs)

    def size(self):
        return len(self.image_names)

    def num_classes(self):
        return max(self.classes.values()) + 1

    def name_to_label(self, name):
        return self.classes[name]

    def label_to_name(self, label):
        return self.labels[label]

    def image_path(self, image_index):
        return os.path.join(self.base_dir, self.image_names[image_index])

    def image_aspect_ratio(self, image_index):
        with Image.open(self.image_path(image_index)) as im:
            return float(im.width) / float(im.height)

    def load_image(self, image_index):
        return read_image_bgr(self.image_path(image_index))

    def load_annotations(self, image_index):
        annots = self.image_data[self.image_names[image_index]]
        boxes = np.zeros((len(annots), 5))
        for idx, a in enumerate(annots):
            boxes[idx] = [float(a["x1"]), float(a["y1"]), float(a["x2"]), float(a["y2"]), self.name_to_label(a["class"])]
        return boxes
