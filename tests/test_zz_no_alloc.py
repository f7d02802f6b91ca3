"""The allocation-free NMS entry point (``b200_nms_sorted_noalloc``) against the regular one.

Some users of libb200ops.so are not allowed to allocate device memory at call time: a serving
process whose whole memory budget is owned by another component, a plugin host that forbids
cudaMalloc, or a latency-critical loop where an allocation could stall the device.  For them the
library keeps the NMS suppression-mask scratch in a built-in 2 MiB device array (limit: 4096
boxes).  This test drives that entry point through the plain C ABI (ctypes, raw pointers) and checks
that it keeps exactly the boxes the regular entry point keeps -- on two calls in a row, the second
with fewer boxes, so that whatever the first call left in the shared scratch is shown not to
matter -- and that it refuses inputs above its limit.

The file name sorts last on purpose: if this test fails under ``pytest -x`` every other test has
already run.  It does not depend on them: it also passes alone in a fresh process.
"""
import ctypes as C

import pytest
import torch

from batchai_retinanet_horovod_coco_b200.ops import native

pytestmark = pytest.mark.gpu


def _boxes(n, seed):
    g = torch.Generator().manual_seed(seed)
    xy = torch.rand(n, 2, generator=g) * 400
    wh = torch.rand(n, 2, generator=g) * 90 + 4
    boxes = torch.cat([xy, xy + wh], dim=1)
    scores = torch.rand(n, generator=g)
    order = scores.argsort(descending=True)
    cls = torch.randint(0, 3, (n,), generator=g, dtype=torch.int32)
    return boxes[order].contiguous().cuda(), cls[order].contiguous().cuda()


def _regular(lib, boxes, cls, thr, stream):
    n = boxes.shape[0]
    words = (n + 63) // 64
    ws = torch.empty(n * words, dtype=torch.int64, device="cuda")      # caller-provided scratch
    keep = torch.zeros(n, dtype=torch.int32, device="cuda")            # caller must zero it
    rc = lib.b200_nms_sorted(C.c_void_p(boxes.data_ptr()), C.c_void_p(0 if cls is None else cls.data_ptr()), n,
                             C.c_float(thr), C.c_void_p(ws.data_ptr()), C.c_void_p(keep.data_ptr()), stream)
    assert rc == 0, lib.b200_last_error().decode()
    return keep


def _noalloc(lib, boxes, cls, thr, stream):
    n = boxes.shape[0]
    keep = torch.full((n,), 7, dtype=torch.int32, device="cuda")       # garbage: the call clears it itself
    rc = lib.b200_nms_sorted_noalloc(C.c_void_p(boxes.data_ptr()), C.c_void_p(0 if cls is None else cls.data_ptr()),
                                     n, C.c_float(thr), C.c_void_p(keep.data_ptr()), stream)
    assert rc == 0, lib.b200_last_error().decode()
    return keep


def test_noalloc_nms_matches_regular_entry_point_twice():
    lib = native.require()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    limit = int(lib.b200_nms_sorted_noalloc_max())
    assert limit == 4096
    for n, seed, use_cls in ((3000, 0, True), (777, 1, False)):        # two calls in a row, shared scratch
        boxes, cls = _boxes(n, seed)
        cls = cls if use_cls else None
        want = _regular(lib, boxes, cls, 0.5, stream)
        got = _noalloc(lib, boxes, cls, 0.5, stream)
        torch.cuda.synchronize()
        assert torch.equal(got, want), f"n={n}: {(got != want).sum().item()} keep flags differ"
        assert 0 < int(got.sum()) < n


def test_noalloc_nms_rejects_sizes_beyond_its_scratch():
    lib = native.require()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    n = int(lib.b200_nms_sorted_noalloc_max()) + 1
    boxes, _ = _boxes(n, 2)
    keep = torch.zeros(n, dtype=torch.int32, device="cuda")
    rc = lib.b200_nms_sorted_noalloc(C.c_void_p(boxes.data_ptr()), C.c_void_p(0), n, C.c_float(0.5),
                                     C.c_void_p(keep.data_ptr()), stream)
    assert rc != 0
    assert b"exceeds" in lib.b200_last_error()
    torch.cuda.synchronize()
    assert int(keep.sum()) == 0
