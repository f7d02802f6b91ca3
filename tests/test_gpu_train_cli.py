"""train.py on the GPU with the native engine: synthetic data, evaluation, snapshot, resume."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, np_=1, cwd=None, timeout=600):
    env = dict(os.environ, PYTHONPATH=ROOT)
    cmd = [sys.executable, "-m", "batchai_retinanet_horovod_coco_b200.parallel.launcher", "-np", str(np_), sys.executable,
           os.path.join(ROOT, "train.py")] + args
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=cwd, env=env)


def test_native_cli_train_eval_snapshot_resume(tmp_path):
    common = ["--batch-size", "2", "--image-min-side", "192", "--image-max-side", "256", "--snapshot-path",
              str(tmp_path / "snap"), "--tensorboard-dir", str(tmp_path / "logs"), "--lr", "1e-4"]
    r = _run(["--no-weights", "--steps", "4", "--epochs", "1"] + common + ["synthetic", "--synthetic-images", "4",
             "--synthetic-classes", "5", "--synthetic-width", "256"], cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert os.path.exists(tmp_path / "snap" / "checkpoint-01.h5") and "mAP" in r.stdout
    r = _run(["--snapshot", str(tmp_path / "snap" / "checkpoint-01.h5"), "--steps", "2", "--epochs", "2", "--no-evaluation"]
             + common + ["synthetic", "--synthetic-images", "4", "--synthetic-classes", "5", "--synthetic-width", "256"],
             cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "Epoch 2/2" in r.stdout and os.path.exists(tmp_path / "snap" / "checkpoint-02.h5")
    from batchai_retinanet_horovod_coco_b200.utils.keras_h5 import load_full_model
    _, st = load_full_model(str(tmp_path / "snap" / "checkpoint-02.h5"))
    assert st["iterations"] == 6


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_native_cli_two_gpus(tmp_path):
    r = _run(["--no-weights", "--steps", "3", "--epochs", "1", "--no-evaluation", "--batch-size", "1", "--image-min-side",
              "192", "--image-max-side", "256", "--snapshot-path", str(tmp_path / "snap"), "--tensorboard-dir", "",
              "synthetic", "--synthetic-images", "4", "--synthetic-classes", "5", "--synthetic-width", "256"], np_=2,
             cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert os.listdir(tmp_path / "snap") == ["checkpoint-01.h5"]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_native_cli_multi_gpu_towers(tmp_path):
    """--multi-gpu 2 --multi-gpu-force (reference train.py:367-368): one process, two towers, evaluation + snapshot."""
    env = dict(os.environ, PYTHONPATH=ROOT)
    cmd = [sys.executable, os.path.join(ROOT, "train.py"), "--no-weights", "--multi-gpu", "2", "--multi-gpu-force",
           "--batch-size", "2", "--steps", "3", "--epochs", "1", "--image-min-side", "192", "--image-max-side", "256",
           "--snapshot-path", str(tmp_path / "snap"), "--tensorboard-dir", "", "--lr", "1e-4", "synthetic",
           "--synthetic-images", "4", "--synthetic-classes", "5", "--synthetic-width", "256"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=str(tmp_path), env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert os.listdir(tmp_path / "snap") == ["checkpoint-01.h5"] and "mAP" in r.stdout
