"""ctypes binding of ``libb200rt.so``'s collective runtime (``csrc/rt/runtime.cpp``) and the
process-level state behind the Horovod-style API in ``parallel/__init__.py``.

Rank discovery accepts, in order: our launcher (``B200_RANK`` ...), torchrun (``RANK`` ...),
OpenMPI's ``mpirun`` (``OMPI_COMM_WORLD_RANK`` ...; the reference launches with mpirun,
``training_job.json:7``), else a single process.
"""

from __future__ import annotations

import ctypes as C
import os

import torch

_LIB = None
# B200RT_LIB points at an alternative build of the runtime (the ThreadSanitizer build of scripts/tsan_runtime.sh)
_LIB_PATH = os.environ.get("B200RT_LIB") or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_native",
                                                        "libb200rt.so")

ALLREDUCE, BROADCAST, ALLGATHER = 0, 1, 2
_DTYPES = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3, torch.uint8: 4}


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(f"native runtime missing: {_LIB_PATH}; run `python build_native.py`")
        l = C.CDLL(_LIB_PATH)
        l.b200rt_last_error.restype = C.c_char_p
        l.b200rt_enqueue.restype = C.c_longlong
        l.b200rt_enqueue.argtypes = [C.c_int, C.c_char_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int,
                                     C.c_double]
        l.b200rt_init.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_double, C.c_longlong]
        l.b200rt_poll.argtypes = [C.c_longlong]
        l.b200rt_wait.argtypes = [C.c_longlong]
        l.b200rt_release.argtypes = [C.c_longlong]
        l.b200rt_result_count.argtypes = [C.c_longlong]
        l.b200rt_result_count.restype = C.c_longlong
        l.b200rt_timeline_start.argtypes = [C.c_char_p]
        _LIB = l
    return _LIB


def err() -> str:
    return lib().b200rt_last_error().decode()


def discover_env():
    """(rank, size, local_rank, local_size, job_id) from the environment."""
    e = os.environ
    for keys in (("B200_RANK", "B200_SIZE", "B200_LOCAL_RANK", "B200_LOCAL_SIZE"),
                 ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE"),
                 ("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK",
                  "OMPI_COMM_WORLD_LOCAL_SIZE"),
                 ("PMI_RANK", "PMI_SIZE", "MPI_LOCALRANKID", "MPI_LOCALNRANKS")):
        if keys[0] in e and keys[1] in e:
            rank, size = int(e[keys[0]]), int(e[keys[1]])
            lrank = int(e.get(keys[2], rank))
            lsize = int(e.get(keys[3], size))
            break
    else:
        rank, size, lrank, lsize = 0, 1, 0, 1
    job = e.get("B200_JOB_ID") or e.get("TORCHELASTIC_RUN_ID", "") + "_" + e.get("MASTER_PORT", "") \
        or e.get("OMPI_MCA_orte_hnp_uri", "")
    job = "".join(ch if ch.isalnum() else "_" for ch in (job or f"solo{os.getpid()}"))[:60]
    return rank, size, lrank, lsize, job


class Handle:
    """An in-flight collective.  Keeps the tensors alive until ``wait()``."""

    def __init__(self, h, keep, finalize=None):
        self.h, self.keep, self.finalize = h, keep, finalize
        self.done = False
        self.result = None

    def poll(self) -> bool:
        if self.done:
            return True
        s = lib().b200rt_poll(self.h)
        if s < 0:
            raise RuntimeError(f"collective failed: {err()}")
        return s == 1

    def wait(self):
        if self.done:
            return self.result
        rc = lib().b200rt_wait(self.h)
        msg = err() if rc != 0 else ""
        if rc == 0 and self.finalize is not None:
            self.result = self.finalize(self)
        lib().b200rt_release(self.h)
        self.done = True
        self.keep = None
        if rc != 0:
            raise RuntimeError(f"collective failed: {msg}")
        return self.result


def enqueue(op: int, name: str, inp: torch.Tensor, out: torch.Tensor, count: int, root: int = 0,
            postscale: float = 1.0, finalize=None) -> Handle:
    if inp.dtype not in _DTYPES:
        raise TypeError(f"unsupported dtype for host collectives: {inp.dtype}")
    h = lib().b200rt_enqueue(op, name.encode(), inp.data_ptr(), out.data_ptr(), count, _DTYPES[inp.dtype], root,
                             postscale)
    if h < 0:
        raise RuntimeError(f"enqueue failed: {err()}")
    return Handle(h, (inp, out), finalize)
