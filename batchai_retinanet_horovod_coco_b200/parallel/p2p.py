"""Peer-to-peer NVLink communicator: symmetric (peer-mapped) gradient bucket + fused allreduce.

Replaces Horovod's NCCL/MPI data plane, fusion-buffer memcpys and the separate ``tf.div``
(reference ``train.py:103``).  Each rank ``cudaMalloc``s ONE region

    [ flag pad | 256 B scalars (this rank's clip factor) | fp32 bucket | fp32 result ]

exports it with CUDA IPC, and maps every peer's region.  The engine's gradients are written
straight into the bucket by the wgrad kernels, so there is nothing to copy in or out; the kernel
in ``csrc/allreduce.cu`` reads all peers' buckets over NVLink, applies each peer's clip factor and
the 1/N scale in registers and writes the averaged gradient to a local buffer.

``torch.distributed`` (any backend) is used only as the control plane: to exchange the 64-byte
IPC handles once, and for the rendezvous itself.
"""

from __future__ import annotations

import ctypes as C

import torch

from ..ops import native

SCALARS_BYTES = 256


class P2PComm:
    def __init__(self, rank: int, world: int, device: torch.device, all_gather_bytes):
        """``all_gather_bytes(b: bytes) -> list[bytes]`` is the control-plane exchange."""
        native.require()
        if world not in (1, 2, 4, 8):
            raise ValueError(f"P2P allreduce kernels are instantiated for 1, 2, 4, 8 ranks, got {world}")
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self._gather = all_gather_bytes
        self.flag_bytes = native.flag_pad_bytes()
        self.epoch = 0
        self.base = 0            # local region
        self.peer_base = []      # region base per rank (own pointer for self)
        self.bucket_elems = 0
        self._opened = []
        self.bucket = None
        self.prescale_local = None

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_torch_distributed(cls, device):
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        rank, world = dist.get_rank(), dist.get_world_size()

        def gather(b: bytes):
            out = [None] * world
            dist.all_gather_object(out, b)
            return out

        return cls(rank, world, device, gather)

    def symmetric_bucket(self, nfloats: int) -> torch.Tensor:
        """Allocate the region (once) and return the local fp32 bucket as a torch tensor."""
        if self.bucket is not None:
            if nfloats > self.bucket_elems:
                raise RuntimeError("symmetric bucket already allocated with a smaller size")
            return self.bucket[:nfloats]
        nfloats = (nfloats + 3) // 4 * 4
        self.bucket_elems = nfloats
        total = self.flag_bytes + SCALARS_BYTES + 8 * nfloats   # bucket + peer-writable result buffer
        with torch.cuda.device(self.device):
            self.base = native.ipc_alloc(total)
            handle = native.ipc_get_handle(self.base) if self.world > 1 else b""
            handles = self._gather(handle) if self.world > 1 else [handle]
            for r in range(self.world):
                if r == self.rank:
                    self.peer_base.append(self.base)
                else:
                    p = native.ipc_open_handle(handles[r])
                    self._opened.append(p)
                    self.peer_base.append(p)
        raw = native.RawCudaBuffer(self.base, total).tensor(self.device)
        self._raw = raw
        off = self.flag_bytes
        self.prescale_local = raw[off:off + 4].view(torch.float32)
        self.prescale_local.fill_(1.0)
        self.bucket = raw[off + SCALARS_BYTES:off + SCALARS_BYTES + 4 * nfloats].view(torch.float32)
        self.result = raw[off + SCALARS_BYTES + 4 * nfloats:off + SCALARS_BYTES + 8 * nfloats].view(torch.float32)
        torch.cuda.synchronize(self.device)
        self.barrier_host()
        return self.bucket

    def barrier_host(self):
        if self.world > 1:
            self._gather(b"x")

    # ------------------------------------------------------------------ collectives
    def _args(self, out: torch.Tensor, offset: int, count: int, postscale: float, has_prescale: bool, algo: int,
              push: bool = False):
        a = native.ArHostArgs()
        for r in range(self.world):
            a.bufs[r] = self.peer_base[r] + self.flag_bytes + SCALARS_BYTES
            a.flags[r] = self.peer_base[r]
            a.prescale[r] = self.peer_base[r] + self.flag_bytes
            a.outs[r] = (self.peer_base[r] + self.flag_bytes + SCALARS_BYTES + 4 * self.bucket_elems) if push else 0
        a.out = out.data_ptr()
        a.rank, a.world = self.rank, self.world
        a.offset, a.count = offset, count
        a.postscale = postscale
        a.epoch = self.epoch
        a.algo = algo
        a.max_blocks = 0
        a.has_prescale = int(has_prescale)
        self.epoch = (self.epoch + 4) & 0x7FFFFFFF
        return a

    def allreduce_avg(self, bucket: torch.Tensor, out: torch.Tensor, n: int, prescale: bool = True, algo: int = 0):
        """out[:n] = (1/world) * sum_r prescale_r * bucket_r[:n]  (bucket must be the symmetric bucket)."""
        assert bucket.data_ptr() == self.bucket.data_ptr(), "allreduce source must be the symmetric bucket"
        n4 = (n + 3) // 4 * 4
        # results can be pushed into every rank's `result` buffer when `out` IS that buffer
        push = out.data_ptr() == self.result.data_ptr()
        a = self._args(out, 0, n4, 1.0 / self.world, prescale, algo, push=push)
        native.allreduce_host(a)

    def allreduce_sum_(self, t: torch.Tensor, average: bool = False):
        """Generic in-place allreduce of an arbitrary fp32 CUDA tensor, staged through the bucket."""
        flat = t.reshape(-1)
        n = flat.numel()
        done = 0
        while done < n:
            m = min(self.bucket_elems, n - done)
            m4 = (m + 3) // 4 * 4
            self.bucket[:m].copy_(flat[done:done + m])
            if m4 > m:
                self.bucket[m:m4].zero_()
            tmp = torch.empty(m4, device=self.device, dtype=torch.float32)
            a = self._args(tmp, 0, m4, (1.0 / self.world) if average else 1.0, False, 0)
            native.allreduce_host(a)
            flat[done:done + m].copy_(tmp[:m])
            done += m
        return t

    def broadcast_(self, t: torch.Tensor, root: int = 0):
        """In-place broadcast of an fp32 CUDA tensor from ``root`` (initial weights / Adam slots, X5)."""
        if self.world == 1:
            return t
        flat = t.reshape(-1)
        n = flat.numel()
        done = 0
        while done < n:
            m = min(self.bucket_elems, n - done)
            m4 = (m + 3) // 4 * 4
            if self.rank == root:
                self.bucket[:m].copy_(flat[done:done + m])
            tmp = torch.empty(m4, device=self.device, dtype=torch.float32)
            a = self._args(tmp, 0, m4, 1.0, False, 0)
            native.broadcast_host(a, root)
            if self.rank != root:
                flat[done:done + m].copy_(tmp[:m])
            done += m
        return t

    def close(self):
        torch.cuda.synchronize(self.device)
        self.barrier_host()
        for p in self._opened:
            try:
                native.ipc_close_handle(p)
            except Exception:
                pass
        self._opened = []
