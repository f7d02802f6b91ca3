"""CLI + driver: argument surface of the reference, generators, callbacks, checkpoints, resume, and the
BASELINE.json config-1 plumbing run (2 ranks on CPU through our launcher)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from batchai_retinanet_horovod_coco_b200.bin import train as T
from batchai_retinanet_horovod_coco_b200.preprocessing.synthetic import write_coco_dataset

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parse_args_defaults_match_reference():
    a = T.parse_args(["coco", "/data/coco"])
    assert (a.backbone, a.batch_size, a.epochs, a.steps) == ("resnet50", 1, 50, 10000)
    assert (a.snapshot_path, a.tensorboard_dir) == ("./snapshots", "./logs")
    assert a.imagenet_weights is True and a.snapshots and a.evaluation
    assert (a.image_min_side, a.image_max_side, a.multi_gpu) == (800, 1333, 0)
    a = T.parse_args(["--no-weights", "--no-evaluation", "--steps", "500", "--freeze-backbone", "coco", "x"])
    assert a.imagenet_weights is False and not a.evaluation and a.steps == 500 and a.freeze_backbone
    a = T.parse_args(["oid", "d", "--labels-filter", "a,b", "--version", "v4"])
    assert a.labels_filter == ["a", "b"]
    a = T.parse_args(["csv", "ann.csv", "cls.csv", "--val-annotations", "v.csv"])
    assert a.val_annotations == "v.csv"


def test_check_args_errors():
    with pytest.raises(ValueError, match="Batch size"):
        T.parse_args(["--multi-gpu", "2", "--batch-size", "1", "coco", "x"])
    with pytest.raises(ValueError, match="experimental"):
        T.parse_args(["--multi-gpu", "2", "--batch-size", "2", "coco", "x"])
    with pytest.raises(ValueError, match="snapshots"):
        T.parse_args(["--multi-gpu", "2", "--batch-size", "2", "--snapshot", "s.h5", "coco", "x"])
    with pytest.raises(SystemExit):
        T.parse_args(["--snapshot", "a", "--weights", "b", "coco", "x"])  # mutually exclusive
    with pytest.warns(UserWarning, match="experimental backbone"):
        T.parse_args(["--backbone", "vgg16", "coco", "x"])


def test_coco_generator_and_compact_batches(tmp_path):
    from batchai_retinanet_horovod_coco_b200.preprocessing.coco import CocoGenerator
    from batchai_retinanet_horovod_coco_b200.utils.transform import random_transform_generator
    root = write_coco_dataset(str(tmp_path / "coco"), num_images=6, image_hw=(96, 128), classes=3)
    g = CocoGenerator(root, "train2017", batch_size=2, image_min_side=96, image_max_side=128,
                      transform_generator=random_transform_generator(flip_x_chance=0.5), seed=1)
    assert g.size() == 6 and g.num_classes() == 3
    assert g.label_to_coco_label(1) == 3 and g.coco_label_to_label(5) == 2 and g.label_to_name(0) == "class_0"
    img, gt, cnt, hw = g.next_compact()
    assert img.dtype == torch.uint8 and img.shape == (2, 96, 128, 3) and gt.shape[2] == 5 and cnt.tolist() == [3, 3]
    x, (reg, lab) = g.next()
    n_anchors = sum(h * w * 9 for h, w in [(12, 16), (6, 8), (3, 4), (2, 2), (1, 1)])
    assert x.shape == (2, 96, 128, 3) and reg.shape == (2, n_anchors, 5) and lab.shape == (2, n_anchors, 4)
    assert abs(float(x.mean())) < 40          # Caffe mean subtracted
    # sharding: disjoint halves
    a = CocoGenerator(root, "train2017", batch_size=1, shard=(0, 2), shuffle_groups=False)
    b = CocoGenerator(root, "train2017", batch_size=1, shard=(1, 2), shuffle_groups=False)
    assert not {i for g_ in a.groups for i in g_} & {i for g_ in b.groups for i in g_}


def test_other_generators(tmp_path):
    import cv2
    from batchai_retinanet_horovod_coco_b200.preprocessing.csv_generator import CSVGenerator
    from batchai_retinanet_horovod_coco_b200.preprocessing.kitti import KittiGenerator
    from batchai_retinanet_horovod_coco_b200.preprocessing.pascal_voc import PascalVocGenerator
    img = np.full((60, 80, 3), 127, np.uint8)
    cv2.imwrite(str(tmp_path / "a.jpg"), img)
    (tmp_path / "ann.csv").write_text("a.jpg,5,6,40,50,cow\na.jpg,1,2,20,30,cat\n")
    (tmp_path / "cls.csv").write_text("cow,0\ncat,1\n")
    g = CSVGenerator(str(tmp_path / "ann.csv"), str(tmp_path / "cls.csv"), batch_size=1, image_min_side=60, image_max_side=80)
    assert g.size() == 1 and g.num_classes() == 2 and g.load_annotations(0).tolist() == [[5, 6, 40, 50, 0], [1, 2, 20, 30, 1]]
    voc = tmp_path / "voc"
    for d in ("ImageSets/Main", "Annotations", "JPEGImages"):
        (voc / d).mkdir(parents=True)
    (voc / "ImageSets/Main/trainval.txt").write_text("img1\n")
    cv2.imwrite(str(voc / "JPEGImages/img1.jpg"), img)
    (voc / "Annotations/img1.xml").write_text(
        "<annotation><object><name>dog</name><truncated>0</truncated><difficult>0</difficult><bndbox><xmin>11</xmin>"
        "<ymin>12</ymin><xmax>41</xmax><ymax>52</ymax></bndbox></object></annotation>")
    g = PascalVocGenerator(str(voc), "trainval", batch_size=1)
    assert g.num_classes() == 20 and g.load_annotations(0).tolist() == [[10, 11, 40, 51, 11]]
    kit = tmp_path / "kitti" / "train"
    (kit / "labels").mkdir(parents=True)
    (kit / "images").mkdir()
    cv2.imwrite(str(kit / "images/000.png"), img)
    (kit / "labels/000.txt").write_text("Car 0.0 0 1.5 10.0 20.0 50.0 55.0 1 1 1 1 1 1 1\n")
    g = KittiGenerator(str(tmp_path / "kitti"), subset="train", batch_size=1)
    assert g.load_annotations(0).tolist() == [[10, 20, 50, 55, 0]] and g.num_classes() == 8


def test_reduce_lr_on_plateau_and_checkpoint_names(tmp_path):
    from batchai_retinanet_horovod_coco_b200.callbacks import ModelCheckpoint, ReduceLROnPlateau

    class Tr:
        lr = 1e-5
        saved = []
        model = None

        def save(self, path, epoch=None, weights_only=False):
            self.saved.append((path, epoch))
    tr = Tr()
    cb = ReduceLROnPlateau(monitor="loss", factor=0.1, patience=2, epsilon=1e-4)
    cb.set_trainer(tr)
    cb.on_train_begin()
    for e, l in enumerate([1.0, 0.9, 0.9, 0.9, 0.9]):
        cb.on_epoch_end(e, {"loss": l})
    assert abs(tr.lr - 1e-6) < 1e-12                       # cut once after 2 epochs without improvement
    ck = ModelCheckpoint(str(tmp_path / "snap" / "checkpoint-{epoch:02d}.h5"))
    ck.set_trainer(tr)
    ck.on_epoch_end(0, {})
    assert tr.saved[0][0].endswith("checkpoint-01.h5") and os.path.isdir(tmp_path / "snap")


def test_keras_adam_matches_formula():
    from batchai_retinanet_horovod_coco_b200.parallel.optimizer import KerasAdam
    p = torch.nn.Parameter(torch.tensor([1.0, -2.0]))
    opt = KerasAdam([p], lr=0.1, clipnorm=0.5)
    p.grad = torch.tensor([3.0, 4.0])                        # norm 5 -> scaled by 0.1
    opt.step()
    g = torch.tensor([0.3, 0.4])
    lr_t = 0.1 * (1 - 0.999) ** 0.5 / (1 - 0.9)
    want = torch.tensor([1.0, -2.0]) - lr_t * (0.1 * g) / ((0.001 * g * g).sqrt() + 1e-7)
    assert torch.allclose(p.detach(), want, atol=1e-6)


@pytest.mark.timeout(900)
def test_two_rank_cpu_training_run(tmp_path):
    """BASELINE.json config 1: tiny COCO-format set, batch 1, 2 ranks on CPU (reference train.py:18)."""
    import build_native
    build_native.build()
    root = write_coco_dataset(str(tmp_path / "coco"), num_images=4, image_hw=(96, 128), classes=3)
    snap, logs, tl = tmp_path / "snapshots", tmp_path / "logs", tmp_path / "timeline.json"
    base = [sys.executable, "-m", "batchai_retinanet_horovod_coco_b200.parallel.launcher", "-np", "2", sys.executable,
            os.path.join(ROOT, "train.py"), "--cpu", "--no-weights", "--batch-size", "1", "--image-min-side", "96",
            "--image-max-side", "128", "--snapshot-path", str(snap), "--tensorboard-dir", str(logs), "--lr", "1e-4"]
    env = dict(os.environ, OMP_NUM_THREADS="4", PYTHONPATH=ROOT)
    r = subprocess.run(base + ["--steps", "2", "--epochs", "2", "--timeline", str(tl), "coco", root], capture_output=True,
                       text=True, timeout=800, cwd=str(tmp_path), env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert sorted(os.listdir(snap)) == ["checkpoint-01.h5", "checkpoint-02.h5"]     # rank 0 only, 1-based
    assert any(f.startswith("events.out.tfevents") for f in os.listdir(logs))
    assert r.stdout.count("Total params") == 1                                       # summary on rank 0 only
    ev = json.loads(tl.read_text())
    assert any(e["name"] == "SHM_ALLREDUCE" for e in ev)
    # resume: optimizer state and epoch counter come back, training continues at epoch 3
    base2 = [a for a in base if a != "--no-weights"]          # --snapshot and --no-weights are mutually exclusive
    r2 = subprocess.run(base2 + ["--steps", "1", "--epochs", "3", "--no-evaluation", "--snapshot", "auto",   # newest = 02
                                "coco", root], capture_output=True, text=True, timeout=800,
                        cwd=str(tmp_path), env=env)
    assert r2.returncode == 0, r2.stdout[-3000:] + r2.stderr[-3000:]
    assert "Epoch 3/3" in r2.stdout and "Epoch 1/3" not in r2.stdout
    assert "checkpoint-03.h5" in os.listdir(snap)
    from batchai_retinanet_horovod_coco_b200.utils.keras_h5 import load_full_model
    _, st = load_full_model(str(snap / "checkpoint-03.h5"))
    assert st["iterations"] == 5 and st["epoch"] == 3


@pytest.mark.timeout(600)
def test_experimental_backbone_cli_run(tmp_path):
    """`--backbone` other than ResNet trains through the same CLI (reference train.py:323-324 only warns)."""
    import build_native
    build_native.build()
    root = write_coco_dataset(str(tmp_path / "coco"), num_images=2, image_hw=(96, 128), classes=3)
    snap = tmp_path / "snapshots"
    cmd = [sys.executable, os.path.join(ROOT, "train.py"), "--cpu", "--backbone", "mobilenet128_0.25", "--batch-size", "1",
           "--image-min-side", "96", "--image-max-side", "128", "--snapshot-path", str(snap), "--steps", "2", "--epochs", "1",
           "--lr", "1e-4", "coco", root]
    env = dict(os.environ, OMP_NUM_THREADS="4", PYTHONPATH=ROOT)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=500, cwd=str(tmp_path), env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "experimental backbone" in r.stderr
    assert "not found locally" in r.stderr          # ImageNet file for this backbone is looked up locally, then skipped
    from batchai_retinanet_horovod_coco_b200.utils.keras_h5 import load_full_model
    m, st = load_full_model(str(snap / "checkpoint-01.h5"))
    assert m.backbone_name == "mobilenet128_0.25" and st["iterations"] == 2
    assert "conv_dw_1/depthwise_kernel:0" in st["slots"] and "conv_pw_13_bn/gamma:0" in st["slots"]


def test_batch_prefetcher_order_coverage_and_errors():
    from batchai_retinanet_horovod_coco_b200.preprocessing.prefetch import BatchPrefetcher
    from batchai_retinanet_horovod_coco_b200.preprocessing.synthetic import SyntheticGenerator
    from batchai_retinanet_horovod_coco_b200.utils.transform import random_transform_generator
    mk = lambda: SyntheticGenerator(num_images=12, image_hw=(64, 96), classes=3, batch_size=2, shuffle_groups=False,
                                    image_min_side=64, image_max_side=96)
    ref = mk()
    want = [ref.next_compact()[1] for _ in range(6)]
    with BatchPrefetcher(mk(), workers=1, max_queue_size=3) as pf:              # one worker: the generator's own order
        got = [pf.next()[1] for _ in range(6)]
    assert all(torch.equal(a, b) for a, b in zip(want, got))
    g = mk()
    g.transform_generator = random_transform_generator(flip_x_chance=0.5)       # shared, not re-entrant: must be locked
    with BatchPrefetcher(g, workers=4, max_queue_size=4) as pf:                 # several workers: every group, any order
        got = [pf.next() for _ in range(12)]
    assert all(b[0].dtype == torch.uint8 and b[0].shape == (2, 64, 96, 3) for b in got)
    with BatchPrefetcher(mk(), workers=0) as pf:                                # synchronous mode
        assert torch.equal(pf.next()[1], want[0])

    class Broken(SyntheticGenerator):
        def load_image(self, i):
            raise OSError("unreadable image")
    with pytest.raises(OSError, match="unreadable"):
        with BatchPrefetcher(Broken(num_images=4, image_hw=(32, 32), classes=2, batch_size=2), workers=2) as pf:
            pf.next()


def test_latest_snapshot_picks_highest_epoch(tmp_path):
    assert T.latest_snapshot(str(tmp_path / "missing")) is None and T.latest_snapshot(str(tmp_path)) is None
    for n in ("checkpoint-02.h5", "checkpoint-10.h5", "checkpoint-9.h5", "checkpoint-xx.h5", "other.h5"):
        (tmp_path / n).write_text("")
    assert T.latest_snapshot(str(tmp_path)) == str(tmp_path / "checkpoint-10.h5")


def test_next_raw_describes_the_same_batch_as_the_host_pipeline(tmp_path):
    """``Generator.next_raw()`` (what the GPU batch assembler consumes: decoded images + augmentation matrices + resize
    scales + transformed boxes) against ``next_compact()`` (pixels produced on the host): replaying the raw description
    with OpenCV on the CPU must give the identical batch, so the two data paths differ only in WHERE the warp and the
    resize run.  (The GPU kernels themselves are compared with OpenCV in tests/test_gpu_ops.py.)"""
    import cv2
    import numpy as np
    from batchai_retinanet_horovod_coco_b200.preprocessing.csv_generator import CSVGenerator
    from batchai_retinanet_horovod_coco_b200.utils.transform import random_transform_generator
    rng = np.random.default_rng(1)
    rows = []
    for i, (h, w) in enumerate([(120, 160), (150, 110), (96, 200), (130, 130)]):
        img = cv2.GaussianBlur(rng.integers(0, 256, (h, w, 3), dtype=np.uint8), (7, 7), 2.0)
        path = str(tmp_path / f"im{i}.png")
        cv2.imwrite(path, img)
        rows.append(f"{path},10,12,{w - 20},{h - 15},thing")
    (tmp_path / "ann.csv").write_text("\n".join(rows) + "\n")
    (tmp_path / "cls.csv").write_text("thing,0\n")

    def make():
        tg = random_transform_generator(min_rotation=-0.1, max_rotation=0.1, min_translation=(-0.1, -0.1),
                                        max_translation=(0.1, 0.1), min_shear=-0.1, max_shear=0.1, min_scaling=(0.9, 0.9),
                                        max_scaling=(1.1, 1.1), flip_x_chance=0.5, flip_y_chance=0.5,
                                        prng=np.random.RandomState(5))
        return CSVGenerator(str(tmp_path / "ann.csv"), str(tmp_path / "cls.csv"), transform_generator=tg, batch_size=2,
                            image_min_side=200, image_max_side=300, shuffle_groups=False, seed=3)

    ga, gb = make(), make()
    for _ in range(2):
        himg, hgt, hcnt, hhw = ga.next_compact()
        raw = gb.next_raw()
        assert torch.equal(hcnt, raw["count"]) and torch.allclose(hgt, raw["gt"], atol=1e-3)
        assert [tuple(x) for x in hhw.tolist()] == [tuple(x) for x in raw["out_hw"]]
        border = cv2.BORDER_REPLICATE if raw["border_replicate"] else cv2.BORDER_CONSTANT
        for i, (img, m, (h, w), (sx, sy)) in enumerate(zip(raw["images"], raw["matrices"], raw["out_hw"], raw["scales"])):
            if m is not None:
                img = cv2.warpAffine(img, np.asarray(m)[:2, :], dsize=(img.shape[1], img.shape[0]), flags=cv2.INTER_LINEAR,
                                     borderMode=border, borderValue=raw["cval"])
            out = cv2.resize(img, None, fx=1.0 / sx, fy=1.0 / sy)
            assert out.shape[:2] == (h, w)
            assert np.array_equal(himg[i, :h, :w].numpy(), out)
