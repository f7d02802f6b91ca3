"""Process launcher: ``python -m batchai_retinanet_horovod_coco_b200.parallel.launcher -np N cmd ...``.

Stands in for ``mpirun`` / ``horovodrun`` (reference ``training_job.json:7``, ``train.py:18``): forks
N local processes with rank / size / local-rank environment (ours, torchrun's and OpenMPI's
variable names, so scripts written for any of them work), a fresh job id for the shared-memory
rendezvous, and 127.0.0.1 as the torch.distributed master.  If any rank exits non-zero the whole
group is terminated (MPI's abort-on-failure behaviour).  mpirun-only flags
(``-mca k v``, ``--allow-run-as-root``, ``--hostfile f``) are accepted and ignored.
"""

from __future__ import annotations

import os
import signal
import socket
import subprocess
import sys
import time
import uuid


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def parse(argv):
    np_, i = 1, 0
    while i < len(argv):
        a = argv[i]
        if a in ("-np", "-n", "--np", "-H") and i + 1 < len(argv):
            if a != "-H":
                np_ = int(argv[i + 1])
            i += 2
        elif a in ("-mca", "--mca") and i + 2 < len(argv):
            i += 3
        elif a in ("--hostfile", "-hostfile", "--bind-to", "-bind-to", "--map-by", "-map-by", "-x") and i + 1 < len(argv):
            i += 2
        elif a in ("--allow-run-as-root", "--oversubscribe"):
            i += 1
        else:
            break
    return np_, argv[i:]


def launch(np_: int, cmd, env_extra=None, timeout=None) -> int:
    job = uuid.uuid4().hex[:12]
    port = _free_port()
    procs = []
    for r in range(np_):
        env = dict(os.environ)
        env.update({"B200_RANK": str(r), "B200_SIZE": str(np_), "B200_LOCAL_RANK": str(r), "B200_LOCAL_SIZE": str(np_),
                    "B200_JOB_ID": job, "RANK": str(r), "WORLD_SIZE": str(np_), "LOCAL_RANK": str(r),
                    "LOCAL_WORLD_SIZE": str(np_), "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
                    "OMPI_COMM_WORLD_RANK": str(r), "OMPI_COMM_WORLD_SIZE": str(np_),
                    "OMPI_COMM_WORLD_LOCAL_RANK": str(r), "OMPI_COMM_WORLD_LOCAL_SIZE": str(np_)})
        env.update(env_extra or {})
        procs.append(subprocess.Popen(cmd, env=env, start_new_session=True))
    t0 = time.time()
    rc = 0
    alive = set(range(np_))
    try:
        while alive:
            for r in list(alive):
                c = procs[r].poll()
                if c is not None:
                    alive.discard(r)
                    if c != 0 and rc == 0:
                        rc = c
                        sys.stderr.write(f"[launcher] rank {r} exited with code {c}; terminating the job\n")
            if rc != 0 or (timeout and time.time() - t0 > timeout):
                if timeout and rc == 0:
                    rc = 124
                break
            time.sleep(0.05)
    finally:
        for r in alive:
            try:
                os.killpg(procs[r].pid, signal.SIGTERM)
            except ProcessLookupError:
                pass
        t1 = time.time()
        for r in alive:
            try:
                procs[r].wait(timeout=max(0.1, 5 - (time.time() - t1)))
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(procs[r].pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass
    return rc


def main(argv=None):
    np_, cmd = parse(list(sys.argv[1:] if argv is None else argv))
    if not cmd:
        sys.stderr.write(__doc__)
        return 2
    return launch(np_, cmd)


if __name__ == "__main__":
    sys.exit(main())
