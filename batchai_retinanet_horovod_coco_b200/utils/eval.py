"""Detection evaluation: VOC-style per-class AP (keras-retinanet ``utils/eval.py``, used by the
``Evaluate`` callback at reference ``train.py:140``) and COCO bbox evaluation
(``utils/coco_eval.py`` / pycocotools, reference ``train.py:135-138``) backed by
the native ``b200rt_cocoeval_bbox`` and ``b200rt_compute_overlap``.

``predictor(images_u8[1,H,W,3]) -> (boxes[1,N,4], scores[1,N], labels[1,N])`` is any callable:
the native engine's ``predict`` or the spec prediction graph.
"""

from __future__ import annotations

import ctypes as C
import json

import numpy as np
import torch

from ..parallel import runtime as _rt
from .image import resize_image


def compute_overlap(boxes: np.ndarray, query: np.ndarray, plus_one: bool = True) -> np.ndarray:
    """IoU matrix (N,K) float64 -- the C++ port of the Cython ``compute_overlap``."""
    boxes = np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 4)
    query = np.ascontiguousarray(query, dtype=np.float64).reshape(-1, 4)
    out = np.zeros((boxes.shape[0], query.shape[0]), dtype=np.float64)
    if out.size:
        _rt.lib().b200rt_compute_overlap(boxes.ctypes.data_as(C.c_void_p), C.c_longlong(boxes.shape[0]),
                                         query.ctypes.data_as(C.c_void_p), C.c_longlong(query.shape[0]),
                                         out.ctypes.data_as(C.c_void_p), int(plus_one))
    return out


def _compute_ap(recall, precision):
    mrec = np.concatenate(([0.0], recall, [1.0]))
    mpre = np.concatenate(([0.0], precision, [0.0]))
    for i in range(mpre.size - 1, 0, -1):
        mpre[i - 1] = np.maximum(mpre[i - 1], mpre[i])
    i = np.where(mrec[1:] != mrec[:-1])[0]
    return np.sum((mrec[i + 1] - mrec[i]) * mpre[i + 1])


def _detect(generator, predictor, index, score_threshold, max_detections):
    raw = generator.load_image(index)
    image, scale = resize_image(raw, generator.image_min_side, generator.image_max_side)
    batch = torch.from_numpy(np.ascontiguousarray(image[None]))
    boxes, scores, labels = predictor(batch)
    boxes = boxes[0].detach().float().cpu().numpy() / scale
    scores = scores[0].detach().float().cpu().numpy()
    labels = labels[0].detach().cpu().numpy()
    keep = np.where(scores > score_threshold)[0]
    order = np.argsort(-scores[keep])[:max_detections]
    keep = keep[order]
    return boxes[keep], scores[keep], labels[keep]


def get_detections(generator, predictor, score_threshold=0.05, max_detections=100, indices=None):
    indices = range(generator.size()) if indices is None else indices
    out = {}
    for i in indices:
        b, s, l = _detect(generator, predictor, i, score_threshold, max_detections)
        out[i] = [np.concatenate([b[l == c], s[l == c, None]], axis=1) for c in range(generator.num_classes())]
    return out


def evaluate(generator, predictor, iou_threshold=0.5, score_threshold=0.05, max_detections=100, indices=None):
    """{label: (average precision, number of annotations)} at ``iou_threshold`` (VOC style)."""
    indices = list(range(generator.size()) if indices is None else indices)
    dets = get_detections(generator, predictor, score_threshold, max_detections, indices)
    anns = {}
    for i in indices:
        a = np.asarray(generator.load_annotations(i)).reshape(-1, 5)
        anns[i] = [a[a[:, 4] == c, :4] for c in range(generator.num_classes())]
    result = {}
    for label in range(generator.num_classes()):
        fp, tp, scores, n_ann = [], [], [], 0
        for i in indices:
            d, a = dets[i][label], anns[i][label]
            n_ann += a.shape[0]
            detected = set()
            ov_all = compute_overlap(d[:, :4], a) if a.shape[0] and d.shape[0] else None
            for k in range(d.shape[0]):
                scores.append(d[k, 4])
                if ov_all is None:
                    fp.append(1); tp.append(0)
                    continue
                j = int(np.argmax(ov_all[k]))
                if ov_all[k, j] >= iou_threshold and j not in detected:
                    fp.append(0); tp.append(1); detected.add(j)
                else:
                    fp.append(1); tp.append(0)
        if n_ann == 0:
            result[label] = (0.0, 0)
            continue
        order = np.argsort(-np.asarray(scores))
        fpc, tpc = np.cumsum(np.asarray(fp)[order]), np.cumsum(np.asarray(tp)[order])
        recall = tpc / n_ann
        precision = tpc / np.maximum(tpc + fpc, np.finfo(np.float64).eps)
        result[label] = (float(_compute_ap(recall, precision)), n_ann)
    return result


def mean_ap(aps: dict) -> float:
    present = [ap for ap, n in aps.values() if n > 0]
    return float(np.mean(present)) if present else 0.0


COCO_STAT_NAMES = ["AP @[ IoU=0.50:0.95 | area=   all | maxDets=100 ]", "AP @[ IoU=0.50      | area=   all | maxDets=100 ]",
                   "AP @[ IoU=0.75      | area=   all | maxDets=100 ]", "AP @[ IoU=0.50:0.95 | area= small | maxDets=100 ]",
                   "AP @[ IoU=0.50:0.95 | area=medium | maxDets=100 ]", "AP @[ IoU=0.50:0.95 | area= large | maxDets=100 ]",
                   "AR @[ IoU=0.50:0.95 | area=   all | maxDets=  1 ]", "AR @[ IoU=0.50:0.95 | area=   all | maxDets= 10 ]",
                   "AR @[ IoU=0.50:0.95 | area=   all | maxDets=100 ]", "AR @[ IoU=0.50:0.95 | area= small | maxDets=100 ]",
                   "AR @[ IoU=0.50:0.95 | area=medium | maxDets=100 ]", "AR @[ IoU=0.50:0.95 | area= large | maxDets=100 ]"]


def cocoeval_bbox(gt_rows, dt_rows, img_ids, cat_ids, gt_area=None) -> np.ndarray:
    """gt_rows [ng,7] (image_id, category_id, x, y, w, h, iscrowd), dt_rows [nd,7] (..., score)."""
    gt = np.ascontiguousarray(gt_rows, dtype=np.float64).reshape(-1, 7)
    dt = np.ascontiguousarray(dt_rows, dtype=np.float64).reshape(-1, 7)
    ids = np.ascontiguousarray(img_ids, dtype=np.int64)
    cats = np.ascontiguousarray(cat_ids, dtype=np.int64)
    area = None if gt_area is None else np.ascontiguousarray(gt_area, dtype=np.float64)
    stats = np.zeros(12, dtype=np.float64)
    _rt.lib().b200rt_cocoeval_bbox(gt.ctypes.data_as(C.c_void_p), None if area is None else area.ctypes.data_as(C.c_void_p),
                                   C.c_longlong(gt.shape[0]), dt.ctypes.data_as(C.c_void_p), C.c_longlong(dt.shape[0]),
                                   ids.ctypes.data_as(C.c_void_p), C.c_longlong(ids.size),
                                   cats.ctypes.data_as(C.c_void_p), C.c_longlong(cats.size),
                                   stats.ctypes.data_as(C.c_void_p))
    return stats


def evaluate_coco(generator, predictor, threshold=0.05, results_path=None, indices=None, verbose=True):
    """COCO bbox evaluation of a ``CocoGenerator``.  Writes ``<set>_bbox_results.json`` like upstream
    (to ``results_path``; pass a per-rank path when several ranks evaluate).  Returns 12 stats or None."""
    indices = list(range(generator.size()) if indices is None else indices)
    results, image_ids = [], []
    for i in indices:
        b, s, l = _detect(generator, predictor, i, threshold, 100)
        image_id = generator.image_ids[i]
        image_ids.append(image_id)
        for box, score, label in zip(b, s, l):
            results.append({"image_id": image_id, "category_id": generator.label_to_coco_label(int(label)),
                            "score": float(score), "bbox": [float(box[0]), float(box[1]), float(box[2] - box[0]),
                                                            float(box[3] - box[1])]})
    if not results:
        return None
    path = results_path or f"{generator.set_name}_bbox_results.json"
    with open(path, "w") as f:
        json.dump(results, f, indent=4)
    wanted = set(image_ids)
    gt_rows, gt_area = [], []
    for a in generator.coco.get("annotations", []):
        if a["image_id"] in wanted:
            gt_rows.append([a["image_id"], a["category_id"], *a["bbox"], float(a.get("iscrowd", 0))])
            gt_area.append(float(a.get("area", a["bbox"][2] * a["bbox"][3])))
    dt_rows = [[r["image_id"], r["category_id"], *r["bbox"], r["score"]] for r in results]
    cat_ids = sorted(c["id"] for c in generator.coco["categories"])
    stats = cocoeval_bbox(gt_rows, dt_rows, image_ids, cat_ids, gt_area)
    if verbose:
        for name, v in zip(COCO_STAT_NAMES, stats):
            kind = "Average Precision" if name.startswith("AP") else "Average Recall"
            print(f" {kind:<18} ({name[:2]}) {name[3:]} = {v:0.3f}")
    return stats
