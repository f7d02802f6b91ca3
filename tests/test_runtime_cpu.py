"""Host collective runtime (libb200rt.so) with real processes on CPU: negotiation, fusion,
allreduce/broadcast/allgather, error responses, stall warning, job abort, timeline, launcher."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "workers", "rt_collectives.py")


def _launch(np_, args, env=None, timeout=120):
    e = dict(os.environ)
    e.update(env or {})
    cmd = [sys.executable, "-m", "batchai_retinanet_horovod_coco_b200.parallel.launcher", "-np", str(np_),
           sys.executable, WORKER] + args
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT, env=e)


@pytest.fixture(scope="module", autouse=True)
def _built():
    import build_native
    build_native.build()


@pytest.mark.parametrize("np_", [2, 3])
def test_collectives_and_fusion(np_, tmp_path):
    tl = tmp_path / "timeline.json"
    r = _launch(np_, ["basic", str(tl)])
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("WORKER_OK") == np_
    ev = json.loads(tl.read_text())
    phases = {e["name"] for e in ev}
    assert {"NEGOTIATE_ALLREDUCE", "MEMCPY_IN_FUSION_BUFFER", "SHM_ALLREDUCE", "MEMCPY_OUT_FUSION_BUFFER"} <= phases


def test_small_fusion_buffer_chunks_large_tensors():
    r = _launch(2, ["basic"], env={"B200RT_FUSION_THRESHOLD": str(1 << 20)})
    assert r.returncode == 0, r.stdout + r.stderr


def test_allreduce_followed_by_broadcast_in_one_response():
    r = _launch(3, ["ar_then_bcast"], env={"B200RT_FAULT": "delay:1:3"})
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("WORKER_OK") == 3


def test_shape_mismatch_is_reported_on_every_rank():
    r = _launch(2, ["mismatch"])
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("MISMATCH_OK") == 2


def test_stall_warning_names_missing_ranks():
    r = _launch(2, ["stall"], env={"B200RT_STALL_CHECK_TIME": "0.5"})
    assert r.returncode == 0, r.stdout + r.stderr
    assert "waiting" in r.stderr and "ranks [1]" in r.stderr


def test_rank_death_terminates_the_job():
    r = _launch(2, ["die"], timeout=60)
    assert r.returncode == 5
    assert "terminating the job" in r.stderr


def test_launcher_accepts_mpirun_style_flags():
    from batchai_retinanet_horovod_coco_b200.parallel.launcher import parse
    np_, cmd = parse(["-mca", "btl_tcp_if_exclude", "docker0,lo", "--allow-run-as-root", "--hostfile", "h", "-np", "8",
                      "python", "train.py", "--steps", "500"])
    assert np_ == 8 and cmd == ["python", "train.py", "--steps", "500"]


def test_launcher_job_spec(tmp_path):
    """--job: the reference's training_job.json layout (nodeCount / processCount / commandLine / output dirs / logs)."""
    import json, subprocess, sys, os
    from batchai_retinanet_horovod_coco_b200.parallel import launcher
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = {"properties": {"nodeCount": 2, "customToolkitSettings": {"processCount": 3, "commandLine":
            "mpirun -mca btl_tcp_if_exclude docker0,lo --allow-run-as-root --hostfile $AZ_BATCHAI_MPI_HOST_FILE "
            + sys.executable + " -c \"import os;open(os.environ['AZ_BATCHAI_OUTPUT_OUTPUTFILES']+'/r'+os.environ['RANK'],'w')"
            ".write(os.environ['WORLD_SIZE']);print('hello from',os.environ['RANK'])\""},
            "stdOutErrPathPrefix": "$AZ_BATCHAI_MOUNT_ROOT/logs",
            "outputDirectories": [{"id": "OUTPUTFILES", "pathPrefix": "$AZ_BATCHAI_MOUNT_ROOT/afs", "pathSuffix": "output"}]}}
    path = tmp_path / "job.json"
    path.write_text(json.dumps(spec))
    np_, argv, env, log_dir = launcher.load_job(str(path))
    assert np_ == 3 and argv[0] == sys.executable and argv[1] == "-c" and log_dir == str(tmp_path / "logs")
    r = subprocess.run([sys.executable, "-m", "batchai_retinanet_horovod_coco_b200.parallel.launcher", "--job", str(path)],
                       capture_output=True, text=True, timeout=120, env=dict(os.environ, PYTHONPATH=root))
    assert r.returncode == 0, r.stderr
    out = tmp_path / "afs" / "output"
    assert sorted(os.listdir(out)) == ["r0", "r1", "r2"] and (out / "r1").read_text() == "3"
    assert "hello from 0" in r.stdout and "hello from 2" in (tmp_path / "logs" / "stdout-2.txt").read_text()
    # the shipped spec parses and targets 8 ranks on one node
    np8, argv8, env8, _ = launcher.load_job(os.path.join(root, "configs", "training_job.json"), mount_root=str(tmp_path))
    assert np8 == 8 and argv8[:2] == ["python", "train.py"] and argv8[-1] == str(tmp_path / "coco")
