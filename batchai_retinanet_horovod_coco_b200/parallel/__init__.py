"""Horovod-equivalent distributed runtime (``import batchai_retinanet_horovod_coco_b200.parallel as hvd``).

Mirrors the ``horovod.keras`` surface the reference binds (``train.py:20-21,71,103,111-112,119``;
SURVEY.md C1-C12): ``init / rank / size / local_rank / local_size``, ``allreduce``, ``allgather``,
``broadcast``, ``broadcast_parameters`` (= BroadcastGlobalVariablesCallback), ``DistributedOptimizer``,
``Compression``, plus timeline control.

Data planes
-----------
* host tensors  -> ``libb200rt.so``: background cycle thread, negotiation by name, tensor fusion,
  shared-memory reduce-scatter/all-gather (the ``mpiexec -n 2`` CPU plumbing config);
* CUDA tensors  -> ``parallel.p2p.P2PComm``: hand-written NVLink peer-to-peer kernels; the IPC
  handles are exchanged through the host runtime's allgather, so no NCCL/MPI is involved.
The native training engine does not go through this dynamic path at all: its gradients live in a
static peer-mapped bucket (see ``engine/trainer.py``).
"""

from __future__ import annotations

import atexit
import io
import os
import pickle

import torch

from . import runtime as _rt

_state = {"init": False, "rank": 0, "size": 1, "local_rank": 0, "local_size": 1, "p2p": None, "counter": 0}


class Compression:
    """``hvd.Compression``: optional fp16 wire format for host allreduce."""

    class none:
        @staticmethod
        def compress(t):
            return t, None

        @staticmethod
        def decompress(t, ctx):
            return t

    class fp16:
        """Half-width wire.  CUDA tensors: the peers' slices really cross NVLink in 16 bits (bf16, same width as
        Horovod's fp16, wider exponent: ``P2PComm.allreduce_range(half_wire=True)``); the reduction stays fp32.
        Host tensors travel through shared memory on one node, where there is no wire to narrow: values are rounded
        to fp16 and back so that results match a compressed run."""
        half_wire = True

        @staticmethod
        def compress(t):
            if t.is_cuda:
                return t, None
            if t.dtype in (torch.float32, torch.float64):
                return t.to(torch.float16).to(torch.float32), t.dtype
            return t, None

        @staticmethod
        def decompress(t, ctx):
            return t if ctx is None else t.to(ctx)


def init(comm=None):
    """Initialise the runtime (idempotent).  Must run before any GPU work is pinned, like
    ``hvd.init()`` at reference ``train.py:21``."""
    if _state["init"]:
        return
    rank, size, lrank, lsize, job = _rt.discover_env()
    cycle = float(os.environ.get("B200RT_CYCLE_TIME", os.environ.get("HOROVOD_CYCLE_TIME", "5")))
    fusion = int(os.environ.get("B200RT_FUSION_THRESHOLD", os.environ.get("HOROVOD_FUSION_THRESHOLD", str(64 << 20))))
    rc = _rt.lib().b200rt_init(rank, size, lrank, lsize, job.encode(), cycle, fusion)
    if rc != 0:
        raise RuntimeError(f"b200rt_init failed: {_rt.err()}")
    _state.update(init=True, rank=rank, size=size, local_rank=lrank, local_size=lsize)
    tl = os.environ.get("B200RT_TIMELINE") or os.environ.get("HOROVOD_TIMELINE")
    if tl and rank == 0:
        _rt.lib().b200rt_timeline_start(tl.encode())
    atexit.register(shutdown)


def shutdown():
    if not _state["init"]:
        return
    if _state["p2p"] is not None:
        try:
            _state["p2p"].close()
        except Exception:
            pass
        _state["p2p"] = None
    _rt.lib().b200rt_shutdown()
    _state["init"] = False


def is_initialized() -> bool:
    return _state["init"]


def _need():
    if not _state["init"]:
        raise ValueError("Horovod-style runtime has not been initialized; use parallel.init().")


def rank() -> int:
    _need()
    return _state["rank"]


def size() -> int:
    _need()
    return _state["size"]


def local_rank() -> int:
    _need()
    return _state["local_rank"]


def local_size() -> int:
    _need()
    return _state["local_size"]


def _auto_name(prefix):
    _state["counter"] += 1
    return f"{prefix}.noname.{_state['counter']}"


def timeline_start(path: str):
    _need()
    if _state["rank"] == 0:
        _rt.lib().b200rt_timeline_start(path.encode())


def timeline_stop():
    if _state["init"]:
        _rt.lib().b200rt_timeline_stop()


def timeline_active() -> bool:
    return bool(_state["init"] and _rt.lib().b200rt_timeline_active())


def timeline_now_ns() -> int:
    import ctypes as C
    f = _rt.lib().b200rt_now_ns
    f.restype = C.c_ulonglong
    return int(f())


def timeline_event(phase: str, name: str, t0_ns: int, t1_ns: int):
    """One complete event of the GPU path (device-timed) in the runtime's Chrome trace, next to the host phases."""
    import ctypes as C
    if timeline_active():
        _rt.lib().b200rt_timeline_event(phase.encode(), name.encode(), C.c_ulonglong(int(t0_ns)), C.c_ulonglong(int(t1_ns)))


def stats() -> dict:
    import ctypes as C
    out = (C.c_longlong * 5)()
    _rt.lib().b200rt_stats(out)
    return dict(zip(("cycles", "responses", "fused_tensors", "bytes_reduced", "stall_warnings"), list(out)))


# ---------------------------------------------------------------------------------- GPU plane
def p2p_comm(device=None, bucket_floats: int = 16 << 20):
    """The process-wide P2P communicator (created on first use)."""
    _need()
    if _state["p2p"] is None:
        from .p2p import P2PComm
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        comm = P2PComm(_state["rank"], _state["size"], dev, _allgather_bytes)
        comm.symmetric_bucket(bucket_floats)
        _state["p2p"] = comm
    return _state["p2p"]


def _allgather_bytes(b: bytes):
    n = size()
    ln = torch.tensor([len(b)], dtype=torch.int64)
    lens = torch.zeros(n, dtype=torch.int64)
    _rt.enqueue(_rt.ALLGATHER, _auto_name("agb.len"), ln, lens, 1).wait()
    buf = torch.frombuffer(bytearray(b), dtype=torch.uint8) if len(b) else torch.zeros(0, dtype=torch.uint8)
    out = torch.zeros(int(lens.sum()), dtype=torch.uint8)
    _rt.enqueue(_rt.ALLGATHER, _auto_name("agb.data"), buf, out, len(b)).wait()
    res, off = [], 0
    for l in lens.tolist():
        res.append(bytes(out[off:off + l].numpy().tobytes()))
        off += l
    return res


# ---------------------------------------------------------------------------------- GPU tensor fusion
class _GpuGroup:
    """Tensors fused into one range of the symmetric bucket; one collective, one completion event."""

    def __init__(self):
        self.items = []          # (tensor, offset, numel)
        self.floats = 0
        self.event = None        # recorded on the comm stream once results are back in the tensors
        self.average = True
        self.half = False


class _GpuFusion:
    """Horovod's tensor fusion for CUDA tensors without its memcpy round trip through a host thread:
    ``allreduce_async_`` copies the tensor into the next free range of the peer-mapped bucket ON A SIDE STREAM (after an
    event on the producer's stream), and when the fused range reaches the fusion threshold -- or someone waits -- ONE
    P2P kernel reduces it and the results are copied back, all on the side stream, so collectives overlap whatever
    the caller keeps launching (backward).  Every rank must enqueue the same tensors in the same order (autograd hooks
    of identical models do); there is no negotiation on this path."""

    def __init__(self, comm):
        self.comm = comm
        self.stream = torch.cuda.Stream(device=comm.device)
        self.threshold = int(os.environ.get("HOROVOD_FUSION_THRESHOLD", os.environ.get("B200RT_FUSION_THRESHOLD", 64 << 20))) // 4
        self.threshold = max(1024, min(self.threshold, comm.bucket_elems))
        self.open = None
        self.done_prev = None    # the bucket is one buffer: the next group starts after the previous one finished

    def add(self, tensor, average, half):
        n = tensor.numel()
        n4 = (n + 3) // 4 * 4
        if n4 > self.comm.bucket_elems:
            raise ValueError(f"tensor of {n} elements exceeds the P2P bucket ({self.comm.bucket_elems}); "
                             "create the communicator with a larger bucket_floats")
        g = self.open
        if g is not None and (g.floats + n4 > self.threshold or g.average != average or g.half != half):
            self.flush()
            g = None
        if g is None:
            g = self.open = _GpuGroup()
            g.average, g.half = average, half
            if self.done_prev is not None:
                self.stream.wait_event(self.done_prev)
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(tensor.device))
        self.stream.wait_event(ready)
        with torch.cuda.stream(self.stream):
            self.comm.bucket[g.floats:g.floats + n].copy_(tensor.reshape(-1), non_blocking=True)
            if n4 > n:
                self.comm.bucket[g.floats + n:g.floats + n4].zero_()
        g.items.append((tensor, g.floats, n))
        g.floats += n4
        return g

    def flush(self):
        g, self.open = self.open, None
        if g is None or g.event is not None:
            return
        with torch.cuda.stream(self.stream):
            self.comm.allreduce_range(0, g.floats, g.average, half_wire=g.half)
            for t, off, n in g.items:
                t.reshape(-1).copy_(self.comm.result[off:off + n], non_blocking=True)
            g.event = torch.cuda.Event()
            g.event.record(self.stream)
        self.done_prev = g.event


class _GpuHandle:
    def __init__(self, fusion, group, tensor):
        self.f, self.g, self.result = fusion, group, tensor

    def poll(self) -> bool:
        if self.g.event is None:
            self.f.flush()
        return self.g.event.query()

    def wait(self):
        """Make the result visible to the CURRENT stream (no host sync), like a stream-ordered Horovod handle."""
        if self.g.event is None:
            self.f.flush()
        torch.cuda.current_stream(self.result.device).wait_event(self.g.event)
        return self.result


def _gpu_fusion(device):
    comm = p2p_comm(device)
    if _state.get("gpu_fusion") is None or _state["gpu_fusion"].comm is not comm:
        _state["gpu_fusion"] = _GpuFusion(comm)
    return _state["gpu_fusion"]


# ---------------------------------------------------------------------------------- collectives
def allreduce_async_(tensor: torch.Tensor, average: bool = True, name: str | None = None, compression=None):
    """In-place asynchronous allreduce.  Returns a handle with ``.wait()`` / ``.poll()``.  CUDA fp32 tensors go
    through the fused P2P path (``_GpuFusion``); ``compression=Compression.fp16`` halves their wire."""
    _need()
    name = name or _auto_name("allreduce")
    if tensor.is_cuda:
        if size() == 1:
            return _Done(tensor)
        half = bool(getattr(compression, "half_wire", False))
        if tensor.dtype == torch.float32 and tensor.is_contiguous():
            f = _gpu_fusion(tensor.device)
            return _GpuHandle(f, f.add(tensor, average, half), tensor)
        t32 = tensor.float().contiguous()
        f = _gpu_fusion(tensor.device)
        h = _GpuHandle(f, f.add(t32, average, half), t32)
        h.wait()
        tensor.copy_(t32)
        return _Done(tensor)
    if not tensor.is_contiguous():
        raise ValueError("allreduce_ needs a contiguous tensor")
    post = (1.0 / size()) if average else 1.0
    return _rt.enqueue(_rt.ALLREDUCE, name, tensor, tensor, tensor.numel(), 0, post, finalize=lambda h: tensor)


class _Done:
    def __init__(self, result):
        self.result = result

    def poll(self):
        return True

    def wait(self):
        return self.result


def allreduce_(tensor, average=True, name=None):
    return allreduce_async_(tensor, average, name).wait()


def allreduce(tensor, average=True, name=None, compression=Compression.none):
    t, ctx = compression.compress(tensor.detach().clone().contiguous())
    allreduce_async_(t, average, name, compression=compression).wait()
    return compression.decompress(t, ctx)


def synchronize(handle):
    return handle.wait()


def poll(handle) -> bool:
    return handle.poll()


def allgather(tensor: torch.Tensor, name: str | None = None) -> torch.Tensor:
    """Concatenate along dim 0; first dims may differ across ranks (Horovod semantics)."""
    _need()
    name = name or _auto_name("allgather")
    src = tensor.detach().cpu().contiguous()
    rest = tuple(src.shape[1:])
    inner = 1
    for s in rest:
        inner *= s
    first = torch.tensor([src.shape[0] if src.dim() else 1], dtype=torch.int64)
    firsts = torch.zeros(size(), dtype=torch.int64)
    _rt.enqueue(_rt.ALLGATHER, name + ".shape", first, firsts, 1).wait()
    total = int(firsts.sum())
    out = torch.empty((total,) + rest, dtype=src.dtype)
    _rt.enqueue(_rt.ALLGATHER, name, src, out, src.numel()).wait()
    return out.to(tensor.device)


def broadcast_(tensor: torch.Tensor, root_rank: int = 0, name: str | None = None):
    _need()
    name = name or _auto_name("broadcast")
    if tensor.is_cuda and tensor.dtype == torch.float32 and tensor.is_contiguous():
        p2p_comm(tensor.device).broadcast_(tensor, root_rank)
        return tensor
    host = tensor.detach().cpu().contiguous()
    view = host if host.dtype in _rt._DTYPES else host.view(torch.uint8)
    _rt.enqueue(_rt.BROADCAST, name, view, view, view.numel(), root_rank).wait()
    if host.data_ptr() != tensor.data_ptr():
        tensor.data.copy_(host)
    return tensor


def broadcast(tensor, root_rank=0, name=None):
    out = tensor.detach().clone()
    return broadcast_(out, root_rank, name)


def broadcast_object(obj, root_rank: int = 0, name: str | None = None):
    _need()
    name = name or _auto_name("bcast_obj")
    data = pickle.dumps(obj) if rank() == root_rank else b""
    ln = torch.tensor([len(data)], dtype=torch.int64)
    _rt.enqueue(_rt.BROADCAST, name + ".len", ln, ln, 1, root_rank).wait()
    buf = torch.frombuffer(bytearray(data), dtype=torch.uint8) if rank() == root_rank else \
        torch.zeros(int(ln.item()), dtype=torch.uint8)
    _rt.enqueue(_rt.BROADCAST, name, buf, buf, buf.numel(), root_rank).wait()
    return pickle.load(io.BytesIO(buf.numpy().tobytes()))


def barrier():
    _need()
    if _rt.lib().b200rt_barrier() != 0:
        raise RuntimeError(f"barrier failed: {_rt.err()}")


def broadcast_parameters(params, root_rank: int = 0):
    """Broadcast a ``state_dict`` / ``named_parameters()`` / module from ``root_rank`` -- weights AND
    buffers (BN statistics), like BroadcastGlobalVariablesCallback(0) at reference ``train.py:111``."""
    if isinstance(params, torch.nn.Module):
        params = params.state_dict()
    items = sorted(params.items()) if isinstance(params, dict) else sorted(params)
    for name, p in items:
        if isinstance(p, torch.Tensor):
            broadcast_(p.data if isinstance(p, torch.nn.Parameter) else p, root_rank, name=f"bcast.param.{name}")


def broadcast_optimizer_state(optimizer, root_rank: int = 0):
    """Broadcast optimizer slots and hyper-parameters (Adam m, v, iterations, lr) from ``root_rank``."""
    inner = getattr(optimizer, "_optimizer", optimizer)
    sd = broadcast_object(inner.state_dict() if rank() == root_rank else None, root_rank, name="bcast.opt_state")
    if rank() != root_rank:
        # make sure slots exist with the right device before loading
        inner.load_state_dict(sd)


from .optimizer import DistributedOptimizer, KerasAdam  # noqa: E402,F401
