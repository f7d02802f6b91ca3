"""Native bbox COCOeval, IoU and VOC-AP against hand-computed fixtures (no pycocotools on the box)."""
import numpy as np

from batchai_retinanet_horovod_coco_b200.ops.anchors import compute_overlap as torch_overlap
from batchai_retinanet_horovod_coco_b200.utils import eval as E


def test_compute_overlap_matches_spec():
    import torch
    rng = np.random.RandomState(0)
    a = rng.uniform(0, 100, (50, 2))
    boxes = np.concatenate([a, a + rng.uniform(5, 60, (50, 2))], axis=1)
    b = rng.uniform(0, 100, (7, 2))
    query = np.concatenate([b, b + rng.uniform(5, 60, (7, 2))], axis=1)
    got = E.compute_overlap(boxes, query)
    want = torch_overlap(torch.from_numpy(boxes), torch.from_numpy(query)).numpy()
    assert np.allclose(got, want, atol=1e-12)
    assert np.allclose(E.compute_overlap(boxes, query, plus_one=False),
                       torch_overlap(torch.from_numpy(boxes), torch.from_numpy(query), plus_one=False).numpy(), atol=1e-12)


def _ap101(pr_at_recall):
    """pr_at_recall: list of (max recall threshold inclusive, precision); beyond the last -> 0."""
    tot = 0.0
    for r in range(101):
        thr = r * 0.01
        p = 0.0
        for rmax, prec in pr_at_recall:
            if thr <= rmax + 1e-12:
                p = prec
                break
        tot += p
    return tot / 101


def test_cocoeval_hand_fixture():
    # one image, one category, two medium-size ground-truth boxes
    gt = [[1, 7, 10, 10, 50, 50, 0], [1, 7, 100, 100, 60, 60, 0]]
    # det A: exact match (IoU 1); det B: false positive; det C: IoU with gt 2 = 50*60/(60*60) = 0.8333 -> wait, shrink
    dt = [[1, 7, 10, 10, 50, 50, 0.9], [1, 7, 300, 300, 40, 40, 0.8], [1, 7, 100, 100, 60, 37.5, 0.7]]  # IoU C = 0.625
    s = E.cocoeval_bbox(gt, dt, [1], [7])
    hi = _ap101([(0.5, 1.0), (1.0, 2.0 / 3.0)])     # thresholds .50 .55 .60: A, C match
    lo = _ap101([(0.5, 1.0)])                       # thresholds >= .65: only A matches
    assert abs(s[1] - hi) < 1e-9                    # AP50
    assert abs(s[2] - lo) < 1e-9                    # AP75
    assert abs(s[0] - (3 * hi + 7 * lo) / 10) < 1e-9
    assert s[3] == -1 and abs(s[4] - s[0]) < 1e-12 and s[5] == -1      # both gts are 'medium' (32^2..96^2)
    # AR: maxDets=1 -> only det A counts: recall .5 at every threshold
    assert abs(s[6] - 0.5) < 1e-12
    assert abs(s[8] - (3 * 1.0 + 7 * 0.5) / 10) < 1e-12 and abs(s[7] - s[8]) < 1e-12


def test_cocoeval_crowd_and_ignore():
    # a crowd region absorbs any number of detections without penalising them
    gt = [[1, 1, 0, 0, 200, 200, 1], [1, 1, 300, 300, 50, 50, 0]]
    dt = [[1, 1, 10, 10, 40, 40, 0.9], [1, 1, 60, 60, 40, 40, 0.8], [1, 1, 300, 300, 50, 50, 0.7]]
    s = E.cocoeval_bbox(gt, dt, [1], [1])
    assert abs(s[1] - 1.0) < 1e-9 and abs(s[8] - 1.0) < 1e-9
    # no detections at all -> AP 0 where ground truth exists
    s = E.cocoeval_bbox(gt, np.zeros((0, 7)), [1], [1])
    assert s[0] == 0.0


class _Gen:
    image_min_side, image_max_side = 100, 100

    def __init__(self):
        self.ann = {0: np.array([[10, 10, 50, 50, 0], [60, 60, 90, 90, 1.0]]), 1: np.array([[20, 20, 70, 70, 0.0]])}

    def size(self):
        return 2

    def num_classes(self):
        return 2

    def load_image(self, i):
        return np.zeros((100, 100, 3), np.uint8)

    def load_annotations(self, i):
        return self.ann[i]

    def label_to_name(self, l):
        return str(l)


def test_voc_evaluate_hand_fixture():
    import torch
    g = _Gen()
    calls = {"n": 0}

    def predictor(batch):
        i = calls["n"]
        calls["n"] += 1
        if i == 0:   # image 0: class-0 hit, class-1 miss (wrong place)
            b = [[10, 10, 50, 50], [0, 0, 20, 20]]
            s, l = [0.9, 0.8], [0, 1]
        else:        # image 1: one false positive scored above the hit
            b = [[80, 80, 99, 99], [21, 21, 70, 70]]
            s, l = [0.95, 0.6], [0, 0]
        pad = 300 - len(b)
        return (torch.tensor([b + [[-1] * 4] * pad], dtype=torch.float32), torch.tensor([s + [-1.0] * pad]),
                torch.tensor([l + [-1] * pad]))
    aps = E.evaluate(g, predictor)
    # class 0: sorted dets FP(.95) TP(.9) TP(.6) over 2 gts: precision envelope 2/3 for both recall steps -> AP 2/3
    assert abs(aps[0][0] - 2.0 / 3.0) < 1e-9 and aps[0][1] == 2
    assert aps[1] == (0.0, 1)
    assert abs(E.mean_ap(aps) - 1.0 / 3.0) < 1e-9
