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


def launch(np_: int, cmd, env_extra=None, timeout=None, log_dir=None) -> int:
    """``log_dir``: write each rank's output to ``stdout-<rank>.txt`` / ``stderr-<rank>.txt`` there (rank 0 is
    also echoed to the console), as Batch AI does under ``stdOutErrPathPrefix`` (reference ``training_job.json:9``)."""
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
        out = err = None
        if log_dir and r > 0:
            os.makedirs(log_dir, exist_ok=True)
            out = open(os.path.join(log_dir, f"stdout-{r}.txt"), "w")
            err = open(os.path.join(log_dir, f"stderr-{r}.txt"), "w")
        procs.append(subprocess.Popen(cmd, env=env, start_new_session=True, stdout=out, stderr=err))
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


def load_job(path: str, mount_root: str | None = None):
    """Job spec in the layout of the reference's ``training_job.json``: ``properties.nodeCount``,
    ``customToolkitSettings.processCount`` / ``commandLine``, ``stdOutErrPathPrefix`` and ``outputDirectories``.

    Returns ``(np, argv, env, log_dir)``.  ``$AZ_BATCHAI_MOUNT_ROOT`` resolves to ``mount_root`` (default: the
    directory of the spec), every output directory ``id`` becomes ``$AZ_BATCHAI_OUTPUT_<id>`` and is created, and a
    leading ``mpirun ...`` / ``horovodrun ...`` in the command line is dropped together with its flags (this launcher
    takes its place).  ``nodeCount`` > 1 is folded onto this node with a warning (NVLink does not leave the box).
    """
    import json
    import shlex
    with open(path) as f:
        props = json.load(f).get("properties", {})
    if int(props.get("nodeCount", 1)) != 1:   # the reference's own spec says 2 nodes x 4 GPUs: same 8 ranks, one box
        sys.stderr.write("[launcher] job spec asks for %s nodes; running all %s ranks on this node\n"
                         % (props.get("nodeCount"), tk_count(props)))
    tk = props.get("customToolkitSettings", {})
    env = {"AZ_BATCHAI_MOUNT_ROOT": os.path.abspath(mount_root or os.path.dirname(os.path.abspath(path)))}

    def expand(s: str) -> str:
        return os.path.expandvars(_subst(s, env))

    for d in props.get("outputDirectories", []):
        out = os.path.join(expand(d.get("pathPrefix", ".")), d.get("pathSuffix", ""))
        os.makedirs(out, exist_ok=True)
        env["AZ_BATCHAI_OUTPUT_" + d["id"]] = out
    log_dir = expand(props["stdOutErrPathPrefix"]) if props.get("stdOutErrPathPrefix") else None
    argv = [expand(a) for a in shlex.split(tk.get("commandLine", ""))]
    if argv and os.path.basename(argv[0]) in ("mpirun", "mpiexec", "horovodrun"):
        _, argv = parse(argv[1:])
    return int(tk.get("processCount", 1)), argv, env, log_dir


def tk_count(props) -> int:
    return int(props.get("customToolkitSettings", {}).get("processCount", 1))


def _subst(s: str, env: dict) -> str:
    for k in sorted(env, key=len, reverse=True):
        s = s.replace("${" + k + "}", env[k]).replace("$" + k, env[k])
    return s


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if argv and argv[0] == "--job":
        if len(argv) < 2:
            sys.stderr.write("--job needs a path to a job spec (JSON)\n")
            return 2
        np_, cmd, env, log_dir = load_job(argv[1], mount_root=argv[3] if len(argv) > 3 and argv[2] == "--mount-root" else None)
        return launch(np_, cmd, env_extra=env, log_dir=log_dir)
    np_, cmd = parse(argv)
    if not cmd:
        sys.stderr.write(__doc__)
        return 2
    return launch(np_, cmd)


if __name__ == "__main__":
    sys.exit(main())
