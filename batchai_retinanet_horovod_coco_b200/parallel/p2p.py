"""Peer-to-peer NVLink communicator: symmetric (peer-mapped) gradient bucket + fused allreduce.

Replaces Horovod's NCCL/MPI data plane, fusion-buffer memcpys and the separate ``tf.div``
(reference ``train.py:103``).  Each rank ``cudaMalloc``s ONE region

    [ flag pad | 256 B scalars (this rank's clip factor) | fp32 bucket | fp32 result | staging ]

(staging = world slots of 1/world of the bucket: where the peers' raw slices of the part this rank owns land while
backward is still running, see ``scatter`` / ``allreduce_staged_avg``)

exports it with CUDA IPC, and maps every peer's region.  The engine's gradients are written
straight into the bucket by the wgrad kernels, so there is nothing to copy in or out; the kernel
in ``csrc/allreduce.cu`` reads all peers' buckets over NVLink, applies each peer's clip factor and
the 1/N scale in registers and writes the averaged gradient to a local buffer.

``torch.distributed`` (any backend) is used only as the control plane: to exchange the 64-byte
IPC handles once, and for the rendezvous itself.
"""

from __future__ import annotations

import os
import socket
import threading
import uuid

import torch

from ..ops import native

SCALARS_BYTES = 256


class P2PComm:
    def __init__(self, rank: int, world: int, device: torch.device, all_gather_bytes):
        """``all_gather_bytes(b: bytes) -> list[bytes]`` is the control-plane exchange."""
        native.require()
        if not 1 <= world <= 8:
            raise ValueError(f"P2P allreduce kernels are instantiated for 1..8 ranks (one NVSwitch node), got {world}")
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
        self.external_sync = False   # LocalP2PComm orders ranks with events instead of in-kernel flags
        self.mc_base = 0             # multicast (NVLS) address of the region, 0 when not available
        self.region_bytes = 0
        self._vmm = None             # handles to release on close

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
        self.stage_off = self.flag_bytes + SCALARS_BYTES + 8 * nfloats
        per4 = -(-(nfloats // 4) // self.world)                       # float4 per owner slice
        self.stage_bytes = self.world * per4 * 16
        total = self.stage_off + self.stage_bytes   # bucket + peer-writable result buffer + staging
        with torch.cuda.device(self.device):
            if not self._alloc_region_vmm(total):
                self._alloc_region_ipc(total)
        raw = native.RawCudaBuffer(self.base, total).tensor(self.device)
        self._raw = raw
        off = self.flag_bytes
        self.prescale_local = raw[off:off + 4].view(torch.float32)
        self.prescale_local.fill_(1.0)
        self.err_word = raw[off + 96:off + 104].view(torch.int32)       # {code, peer} of a timed-out flag barrier
        self.err_word.zero_()
        self.bucket = raw[off + SCALARS_BYTES:off + SCALARS_BYTES + 4 * nfloats].view(torch.float32)
        self.result = raw[off + SCALARS_BYTES + 4 * nfloats:off + SCALARS_BYTES + 8 * nfloats].view(torch.float32)
        torch.cuda.synchronize(self.device)
        self.barrier_host()
        return self.bucket

    # ------------------------------------------------------------------ region allocation
    def _alloc_region_ipc(self, total: int):
        """cudaMalloc + CUDA IPC handles (no multicast address)."""
        self.region_bytes = total
        self.base = native.ipc_alloc(total)
        handle = self._export_region(self.base) if self.world > 1 else b""
        handles = self._gather(handle) if self.world > 1 else [handle]
        for r in range(self.world):
            self.peer_base.append(self.base if r == self.rank else self._import_region(handles[r]))

    def _nvls_wanted(self) -> bool:
        return self.world > 1 and not self.external_sync and os.environ.get("B200_P2P_NVLS", "1") != "0"

    def _alloc_region_vmm(self, total: int) -> bool:
        """VMM allocation shared through POSIX fds, every peer mapped unicast, plus an NVSwitch multicast
        mapping of the same memory (``mc_base``) for the NVLS allreduce.  Returns False (after agreeing with
        every rank) when the box does not offer it, in which case the CUDA IPC path is used."""
        if not self._nvls_wanted():
            return False
        dev = self.device.index if self.device.index is not None else torch.cuda.current_device()
        vmm, mc = native.vmm_supported(dev)
        if not all(f == b"1" for f in self._gather(b"1" if (vmm and mc) else b"0")):
            return False
        state = {"maps": [], "handles": []}
        err = None
        size = mem = mc_handle = base = 0
        peer_base = []
        try:
            gran = native.vmm_granularity(dev, self.world)
            size = -(-total // gran) * gran
            mem, mem_fd = native.vmm_create(dev, size)
            state["handles"].append(mem)
            base = native.vmm_map(mem, size, dev, gran)
            state["maps"].append(base)
            native.RawCudaBuffer(base, size).tensor(self.device).zero_()
            torch.cuda.synchronize(self.device)
            fds = [mem_fd]
            if self.rank == 0:
                mc_handle, mc_fd = native.mc_create(self.world, size)
                state["handles"].append(mc_handle)
                fds.append(mc_fd)
            got = self._exchange_fds(fds)               # {peer rank: [fds]}
            peer_base = [0] * self.world
            for r in range(self.world):
                if r == self.rank:
                    peer_base[r] = base
                    continue
                h = native.vmm_import(got[r][0])
                state["handles"].append(h)
                peer_base[r] = native.vmm_map(h, size, dev, gran)
                state["maps"].append(peer_base[r])
            if self.rank != 0:
                mc_handle = native.vmm_import(got[0][1])
                state["handles"].append(mc_handle)
            for fl in list(got.values()) + [fds]:
                for fd in fl:
                    os.close(fd)
            native.mc_add_device(mc_handle, dev)
        except Exception as e:  # noqa: BLE001 - any failure must be agreed on by all ranks before falling back
            err = e
        # every device must be added before anyone binds; a rank that failed makes all ranks fall back
        if not all(f == b"1" for f in self._gather(b"0" if err else b"1")):
            self._release_vmm(state, size)
            if err is not None and os.environ.get("B200_P2P_NVLS") == "1":
                raise err
            return False
        try:
            native.mc_bind(mc_handle, mem, size)
        except Exception as e:  # noqa: BLE001
            err = e
        if not all(f == b"1" for f in self._gather(b"0" if err else b"1")):
            raise RuntimeError(f"NVLS multicast bind failed on some rank ({err}); set B200_P2P_NVLS=0 to use CUDA IPC")
        mc_base = native.vmm_map(mc_handle, size, dev, gran)
        state["maps"].append(mc_base)
        self.base, self.peer_base, self.mc_base, self.region_bytes = base, peer_base, mc_base, size
        self._vmm = (state, size)
        self._gather(b"x")
        return True

    def _release_vmm(self, state, size):
        for ptr in state["maps"]:
            try:
                native.vmm_unmap(ptr, size)
            except Exception:
                pass
        for h in state["handles"]:
            try:
                native.vmm_release(h)
            except Exception:
                pass

    def _exchange_fds(self, fds):
        """Give ``fds`` to every peer and receive theirs (SCM_RIGHTS over abstract Unix sockets; the names travel
        over the control plane).  Returns {peer rank: [fd, ...]} with this process's own copies."""
        name = f"\0b200p2p-{uuid.uuid4().hex}-{self.rank}"
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(name)
        srv.listen(self.world)
        srv.settimeout(60)
        names = [n.decode("latin1") for n in self._gather(name.encode("latin1"))]

        def serve():
            for _ in range(self.world - 1):
                conn, _ = srv.accept()
                with conn:
                    conn.recv(1)
                    socket.send_fds(conn, [b"f"], list(fds))

        t = threading.Thread(target=serve, daemon=True)
        t.start()
        got = {}
        for r in range(self.world):
            if r == self.rank:
                continue
            with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                c.settimeout(60)
                c.connect(names[r])
                c.send(b"r")
                _, rfds, _, _ = socket.recv_fds(c, 16, 4)
                got[r] = list(rfds)
        t.join(timeout=60)
        srv.close()
        return got

    def _before_collective(self):
        pass

    def _after_collective(self):
        pass

    # how a region travels to the peers: CUDA IPC between processes (LocalP2PComm: plain pointers)
    def _export_region(self, base: int) -> bytes:
        return native.ipc_get_handle(base)

    def _import_region(self, blob: bytes) -> int:
        p = native.ipc_open_handle(blob)
        self._opened.append(p)
        return p

    def barrier_host(self):
        if self.world > 1:
            self._gather(b"x")

    # ------------------------------------------------------------------ collectives
    def _args(self, out: torch.Tensor, offset: int, count: int, postscale: float, has_prescale: bool, algo: int,
              push: bool = False, advance: bool = True):
        a = native.ArHostArgs()
        a.wire_bf16 = int(os.environ.get("B200_AR_WIRE", "fp32") == "bf16")
        # flag barriers give up after B200_AR_TIMEOUT_S seconds (default 30; 0 = wait for ever); ~2 GHz SM clock
        a.err = self.base + self.flag_bytes + 96
        a.timeout_cycles = int(float(os.environ.get("B200_AR_TIMEOUT_S", "30")) * 2.0e9)
        for r in range(self.world):
            a.stage[r] = self.peer_base[r] + self.stage_off
            a.bufs[r] = self.peer_base[r] + self.flag_bytes + SCALARS_BYTES
            a.flags[r] = self.peer_base[r]
            a.prescale[r] = self.peer_base[r] + self.flag_bytes
            a.outs[r] = (self.peer_base[r] + self.flag_bytes + SCALARS_BYTES + 4 * self.bucket_elems) if push else 0
        if push and self.mc_base:
            a.mc_buf = self.mc_base + self.flag_bytes + SCALARS_BYTES
            a.mc_out = self.mc_base + self.flag_bytes + SCALARS_BYTES + 4 * self.bucket_elems
        a.out = out.data_ptr()
        a.rank, a.world = self.rank, self.world
        a.offset, a.count = offset, count
        a.postscale = postscale
        a.epoch = self.epoch
        # the flag epoch lives in device memory (scalars + 64) and is bumped on the device after every collective,
        # so a captured CUDA graph of a training step replays with fresh flag values
        a.epoch_ptr = 0 if self.external_sync else self.base + self.flag_bytes + 64
        a.algo = algo
        a.max_blocks = 0
        a.has_prescale = int(has_prescale)
        a.external_sync = int(self.external_sync)
        if advance:
            self.epoch = (self.epoch + 4) & 0xFFFFFFFF    # wraps mod 2^32 like the device-side counter
        return a

    def allreduce_avg(self, bucket: torch.Tensor, out: torch.Tensor, n: int, prescale: bool = True, algo: int = 0):
        """out[:n] = (1/world) * sum_r prescale_r * bucket_r[:n]  (bucket must be the symmetric bucket)."""
        assert bucket.data_ptr() == self.bucket.data_ptr(), "allreduce source must be the symmetric bucket"
        n4 = (n + 3) // 4 * 4
        # results can be pushed into every rank's `result` buffer when `out` IS that buffer
        push = out.data_ptr() == self.result.data_ptr()
        a = self._args(out, 0, n4, 1.0 / self.world, prescale, algo, push=push)
        self._before_collective()
        native.allreduce_host(a)
        self._after_collective()

    # ---- allreduce split around backward: scatter raw slices as they become final, reduce + gather at the end
    def can_stage(self) -> bool:
        """The staged (scatter-during-backward) path.  ``B200_AR_OVERLAP=1`` / ``0`` forces it on / off; by default it
        is used from 4 ranks up: measured at per-GPU batch 16, it is worth 0.05 ms per step on 8 GPUs and costs
        0.15 ms on 2 (one peer: the plain push allreduce is already only ~0.2 ms there)."""
        if self.world < 2 or self.external_sync:
            return False
        env = os.environ.get("B200_AR_OVERLAP")
        return self.world >= 4 if env is None else env != "0"

    def scatter(self, n: int, lo: int, hi: int, stream=None):
        """Send elements [lo, hi) of the bucket (final on this rank, NOT yet clipped) to the staging area of their
        owner rank.  No flags: ordering comes from the barrier at the start of ``allreduce_staged_avg``.  ``n`` must
        be the ``n`` of that call."""
        n4 = (n + 3) // 4 * 4
        a = self._args(self.result, 0, n4, 1.0, False, 0, push=True, advance=False)
        native.ar_scatter(a, lo, min(hi, n4), stream)

    def allreduce_staged_avg(self, n: int, prescale: bool = True):
        """result[:n] = (1/world) * sum_r prescale_r * bucket_r[:n] where every peer slice was already scattered:
        this rank reduces its 1/world slice from LOCAL memory and multicasts (or pushes) it to every rank."""
        n4 = (n + 3) // 4 * 4
        a = self._args(self.result, 0, n4, 1.0 / self.world, prescale, 0, push=True)
        native.ar_reduce_staged(a)

    def allreduce_range(self, offset: int, n: int, average: bool, half_wire: bool = False):
        """result[offset:offset+n] = sum (or mean) over ranks of bucket[offset:offset+n], launched on the current
        stream.  ``half_wire``: the peers' slices cross NVLink as bf16 (scatter + staged reduce), which is what
        ``hvd.Compression.fp16`` selects for CUDA tensors; otherwise the auto-chosen fused kernel."""
        n4 = (n + 3) // 4 * 4
        post = (1.0 / self.world) if average else 1.0
        if half_wire and self.world > 1 and not self.external_sync:
            a = self._args(self.result, offset, n4, post, False, 0, push=True, advance=False)
            a.wire_bf16 = 1
            native.ar_scatter(a, 0, n4)
            a = self._args(self.result, offset, n4, post, False, 0, push=True)
            a.wire_bf16 = 1
            native.ar_reduce_staged(a)
            return
        a = self._args(self.result, offset, n4, post, False, 0, push=True)
        self._before_collective()
        native.allreduce_host(a)
        self._after_collective()

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
            self._before_collective()
            native.allreduce_host(a)
            self._after_collective()
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
            use_mc = bool(self.mc_base)      # through the switch: one multicast store reaches every result buffer
            tmp = self.result if use_mc else torch.empty(m4, device=self.device, dtype=torch.float32)
            a = self._args(tmp, 0, m4, 1.0, False, 4 if use_mc else 0, push=use_mc)
            self._before_collective()
            native.broadcast_host(a, root)
            self._after_collective()
            if self.rank != root:
                flat[done:done + m].copy_(tmp[:m])
            done += m
        return t

    def check(self):
        """Raise if a flag barrier of an earlier collective timed out (a peer did not arrive).  Synchronises."""
        if self.bucket is None or self.world == 1:
            return
        code, peer = (int(v) for v in self.err_word.cpu())
        if code != 0:
            raise RuntimeError(f"P2P collective timed out on rank {self.rank}: rank {peer} never reached the flag "
                               f"barrier (epoch {code & 0x7FFFFFFF}); the job cannot continue")

    def close(self):
        torch.cuda.synchronize(self.device)
        self.barrier_host()
        for p in self._opened:
            try:
                native.ipc_close_handle(p)
            except Exception:
                pass
        self._opened = []
        if self._vmm is not None:
            self.bucket = self.result = self.prescale_local = self._raw = None
            self._release_vmm(*self._vmm)
            self._vmm = None


class LocalGroup:
    """Control plane for ranks that are threads of ONE process (``--multi-gpu`` towers): a host barrier,
    an all-gather of byte strings, and the per-rank CUDA events the ranks order their streams with."""

    TIMEOUT_S = 120.0

    def __init__(self, world: int):
        import threading
        self.world = world
        self._barrier = threading.Barrier(world)
        self._slots = [None] * world
        self.ev_ready = [None] * world
        self.ev_done = [None] * world

    def wait(self):
        self._barrier.wait(timeout=self.TIMEOUT_S)   # BrokenBarrierError if a tower died or is stuck

    def abort(self):
        self._barrier.abort()

    def gather_fn(self, rank: int):
        def gather(b: bytes):
            self._slots[rank] = b
            self.wait()
            out = list(self._slots)
            self.wait()
            return out
        return gather


class LocalP2PComm(P2PComm):
    """Same reduce kernels, ranks = GPUs driven by threads of one process.

    * Regions are shared as raw device pointers after ``cudaDeviceEnablePeerAccess`` (a CUDA IPC handle
      cannot be opened by the process that exported it).
    * No in-kernel flag barriers: inside one process a kernel spinning on a peer can deadlock against an
      implicitly synchronising runtime call (allocation, free) issued for that peer.  The ranks order their
      streams with events instead: record "bucket ready" -> host barrier -> wait on every peer's event ->
      flag-free one-shot reduce -> record "done" -> host barrier -> wait on every peer's "done" before the
      bucket may be overwritten.
    """

    def __init__(self, rank: int, world: int, device, group: LocalGroup):
        super().__init__(rank, world, device, group.gather_fn(rank))
        self.group = group
        self.external_sync = True
        with torch.cuda.device(self.device):
            native.enable_peer_access(self.device.index, world)
            group.ev_ready[rank] = torch.cuda.Event()
            group.ev_done[rank] = torch.cuda.Event()

    def _export_region(self, base: int) -> bytes:
        return int(base).to_bytes(8, "little")

    def _import_region(self, blob: bytes) -> int:
        return int.from_bytes(blob, "little")

    def _sync(self, events):
        stream = torch.cuda.current_stream(self.device)
        events[self.rank].record(stream)
        self.group.wait()                       # every rank has recorded
        for r in range(self.world):
            if r != self.rank:
                stream.wait_event(events[r])
        self.group.wait()                       # every rank has issued its waits: events may be re-recorded

    def _before_collective(self):
        self._sync(self.group.ev_ready)

    def _after_collective(self):
        self._sync(self.group.ev_done)
