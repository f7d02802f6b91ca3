"""Multi-GPU P2P collectives and data-parallel engine step.  Each test spawns one process per GPU
(torchrun-style env, 127.0.0.1 rendezvous) and is skipped when fewer than 2 GPUs are visible."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _run(script, nproc, timeout=420):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", "29617", os.path.join(ROOT, "tests", "multi", script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_p2p_allreduce_and_broadcast():
    n = 8 if _ngpu() >= 8 else (4 if _ngpu() >= 4 else 2)
    out = _run("p2p_collectives.py", n)
    assert "P2P_OK" in out


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_data_parallel_step_matches_single_process():
    out = _run("dp_step.py", 2)
    assert "DP_OK" in out


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_single_process_towers():
    """--multi-gpu 2 (reference train.py:86-89): N engines in one process, same allreduce kernel over peer pointers."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "multi", "towers.py")], capture_output=True, text=True,
                       timeout=420, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "TOWERS_OK" in r.stdout


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_flag_barrier_times_out_when_a_rank_is_missing():
    out = _run("barrier_timeout.py", 2, timeout=180)
    assert "TIMEOUT_OK" in out


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_horovod_style_api_on_cuda_tensors():
    """DistributedOptimizer over torch.optim.SGD, fused async GPU allreduce, Compression.fp16 (reference train.py:103)."""
    cmd = [sys.executable, "-m", "batchai_retinanet_horovod_coco_b200.parallel.launcher", "-np", "2", sys.executable,
           os.path.join(ROOT, "tests", "multi", "hvd_gpu.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("HVD_GPU_OK") == 2
