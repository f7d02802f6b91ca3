"""Pascal VOC generator (keras-retinanet ``preprocessing/pascal_voc.py``; reference ``train.py:216-231``):
``ImageSets/Main/<set>.txt``, ``Annotations/<id>.xml``, ``JPEGImages/<id>.jpg``; 20 classes."""

from __future__ import annotations

import os
import xml.etree.ElementTree as ET

import numpy as np
from PIL import Image

from ..utils.image import read_image_bgr
from .generator import Generator

voc_classes = {"aeroplane": 0, "bicycle": 1, "bird": 2, "boat": 3, "bottle": 4, "bus": 5, "car": 6, "cat": 7,
               "chair": 8, "cow": 9, "diningtable": 10, "dog": 11, "horse": 12, "motorbike": 13, "person": 14,
               "pottedplant": 15, "sheep": 16, "sofa": 17, "train": 18, "tvmonitor": 19}


def _find(parent, name, parse=None):
    node = parent.find(name)
    if node is None:
        raise ValueError(f"missing element '{name}'")
    if parse is None:
        return node
    try:
        return parse(node.text)
    except ValueError as e:
        raise ValueError(f"illegal value for '{name}': {e}") from None


class PascalVocGenerator(Generator):
    def __init__(self, data_dir, set_name, classes=voc_classes, image_extension=".jpg", skip_truncated=False,
                 skip_difficult=False, **kwargs):
        self.data_dir, self.set_name, self.classes = data_dir, set_name, classes
        with open(os.path.join(data_dir, "ImageSets", "Main", set_name + ".txt")) as f:
            self.image_names = [l.strip().split(None, 1)[0] for l in f if l.strip()]
        self.image_extension = image_extension
        self.skip_truncated, self.skip_difficult = skip_truncated, skip_difficult
        self.labels = {v: k for k, v in self.classes.items()}
        super().__init__(**kwargs)

    def size(self):
        return len(self.image_names)

    def num_classes(self):
        return len(self.classes)

    def name_to_label(self, name):
        return self.classes[name]

    def label_to_name(self, label):
        return self.labels[label]

    def image_path(self, image_index):
        return os.path.join(self.data_dir, "JPEGImages", self.image_names[image_index] + self.image_extension)

    def image_aspect_ratio(self, image_index):
        with Image.open(self.image_path(image_index)) as im:
            return float(im.width) / float(im.height)

    def load_image(self, image_index):
        return read_image_bgr(self.image_path(image_index))

    def _parse_annotation(self, element):
        truncated = _find(element, "truncated", int) if element.find("truncated") is not None else 0
        difficult = _find(element, "difficult", int) if element.find("difficult") is not None else 0
        class_name = _find(element, "name").text
        if class_name not in self.classes:
            raise ValueError(f"class name '{class_name}' not found in classes: {list(self.classes)}")
        box = np.zeros((1, 5))
        box[0, 4] = self.name_to_label(class_name)
        bndbox = _find(element, "bndbox")
        box[0, 0] = _find(bndbox, "xmin", float) - 1
        box[0, 1] = _find(bndbox, "ymin", float) - 1
        box[0, 2] = _find(bndbox, "xmax", float) - 1
        box[0, 3] = _find(bndbox, "ymax", float) - 1
        return truncated, difficult, box

    def load_annotations(self, image_index):
        filename = self.image_names[image_index] + ".xml"
        try:
            tree = ET.parse(os.path.join(self.data_dir, "Annotations", filename))
            boxes = np.zeros((0, 5))
            for i, element in enumerate(tree.getroot().iter("object")):
                try:
                    truncated, difficult, box = self._parse_annotation(element)
                except ValueError as e:
                    raise ValueError(f"could not parse object #{i}: {e}") from None
                if (truncated and self.skip_truncated) or (difficult and self.skip_difficult):
                    continue
                boxes = np.append(boxes, box, axis=0)
            return boxes
        except (ET.ParseError, ValueError) as e:
            raise ValueError(f"invalid annotations file: {filename}: {e}") from None
