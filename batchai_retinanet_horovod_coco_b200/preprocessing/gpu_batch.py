"""Batch assembly on the GPU for the real-data path (hard part 9).

The reference's generator decodes, warps (``--random-transform``), resizes to 800/1333 and pads every image on the CPU in
one Python thread (keras-retinanet ``Generator`` as bound at reference ``train.py:177-214``).  Here the host only
DECODES: ``Generator.next_raw()`` hands over the decoded uint8 image, the 3x3 augmentation matrix and the resize scale,
and ``GpuBatchAssembler`` uploads the raw pixels and runs ``b200_warp_affine_u8`` + ``b200_resize_linear_u8`` straight
into the padded batch tensor.  Both kernels use OpenCV's fixed-point arithmetic, so the batch matches the host pipeline
to +-1 LSB (``tests/test_gpu_ops.py``).  Boxes follow the same transform on the host (a few floats per image).
"""

from __future__ import annotations

import threading

import numpy as np
import torch

from ..ops import native


class GpuBatchAssembler:
    """``assemble_async`` is meant to be called from the prefetch WORKER threads: each thread uploads and transforms
    on its own CUDA stream and returns (batch, event); the training thread only makes its stream wait for the event
    (``finish``), so neither the pinned copies nor the kernel launches sit in the step loop."""

    def __init__(self, device, pad_bgr=(104, 117, 124)):
        self.device = torch.device(device)
        self.pad = torch.tensor(pad_bgr, dtype=torch.uint8, device=self.device)
        self._local = threading.local()

    def _state(self):
        st = self._local
        if not hasattr(st, "stream"):
            st.stream = torch.cuda.Stream(device=self.device)
            st.pinned = {}
        return st

    def _stage(self, st, slot: int, img: np.ndarray) -> torch.Tensor:
        n = img.size
        buf = st.pinned.get(slot)
        if buf is None or buf.numel() < n:
            buf = st.pinned[slot] = torch.empty(1 << max(16, int(n - 1).bit_length()), dtype=torch.uint8).pin_memory()
        view = buf[:n].view(img.shape)
        view.copy_(torch.from_numpy(np.ascontiguousarray(img)))
        return view.to(self.device, non_blocking=True)

    def assemble_async(self, raw: dict):
        """raw: {'images': [uint8 HxWx3], 'matrices': [3x3 float64 or None], 'out_hw': [(h, w)], 'scales': [(sx, sy)],
        'border_replicate': bool, 'cval': int, 'gt': (B,G,5), 'count': (B,)}  ->  ((images uint8 [B,H,W,3] on the
        device, gt, count, hw), event recorded on this thread's stream)."""
        st = self._state()
        B = len(raw["images"])
        H = max(h for h, _ in raw["out_hw"])
        W = max(w for _, w in raw["out_hw"])
        with torch.cuda.stream(st.stream):
            st.stream.synchronize()          # this thread's pinned staging buffers are free again
            batch = torch.empty(B, H, W, 3, dtype=torch.uint8, device=self.device)
            batch.view(-1, 3)[:] = self.pad
            for i, img in enumerate(raw["images"]):
                src = self._stage(st, i, img)
                m = raw["matrices"][i]
                if m is not None:
                    warped = torch.empty_like(src)
                    native.warp_affine_u8(src, warped, np.asarray(m)[:2, :],
                                          border_replicate=raw.get("border_replicate", True), cval=raw.get("cval", 0))
                    src = warped
                h, w = raw["out_hw"][i]
                sx, sy = raw["scales"][i]
                native.resize_linear_u8(src, batch[i, :h, :w], sx, sy)
            ev = torch.cuda.Event()
            ev.record(st.stream)
        hw = torch.tensor(raw["out_hw"], dtype=torch.int32)
        return (batch, raw["gt"], raw["count"], hw), ev

    def finish(self, item):
        """Training thread: order the current stream after the worker's and keep the allocator honest."""
        (batch, gt, cnt, hw), ev = item
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        batch.record_stream(cur)
        return batch, gt, cnt, hw

    def assemble(self, raw: dict):
        return self.finish(self.assemble_async(raw))
