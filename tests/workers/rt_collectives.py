import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import batchai_retinanet_horovod_coco_b200.parallel as hvd

hvd.init()
r, n = hvd.rank(), hvd.size()
assert n == int(os.environ["B200_SIZE"]) and hvd.local_rank() == r and hvd.local_size() == n
mode = sys.argv[1] if len(sys.argv) > 1 else "basic"
if mode == "basic":
    if r == 0 and len(sys.argv) > 2:
        hvd.timeline_start(sys.argv[2])
    # many small tensors enqueued together -> fused; one big tensor > fusion threshold -> chunked
    ts = [torch.full((17 + i,), float(r + 1 + i)) for i in range(20)]
    hs = [hvd.allreduce_async_(t, average=True, name=f"g{i}") for i, t in enumerate(ts)]
    for i, h in enumerate(hs):
        h.wait()
        want = sum(q + 1 + i for q in range(n)) / n
        assert torch.allclose(ts[i], torch.full_like(ts[i], want)), (i, ts[i][:3], want)
    big = torch.arange(3_000_000, dtype=torch.float64) * (r + 1)
    hvd.allreduce_(big, average=False, name="big")
    assert torch.equal(big, torch.arange(3_000_000, dtype=torch.float64) * (n * (n + 1) // 2))
    it = torch.tensor([r, 2 * r], dtype=torch.int64)
    assert hvd.allreduce(it, average=False).tolist() == [sum(range(n)), 2 * sum(range(n))]
    b = torch.full((5,), float(r))
    hvd.broadcast_(b, root_rank=n - 1)
    assert (b == n - 1).all()
    g = hvd.allgather(torch.full((r + 1, 3), float(r)))
    assert g.shape == (n * (n + 1) // 2, 3) and g[0, 0] == 0 and g[-1, 0] == n - 1
    obj = hvd.broadcast_object({"lr": 1e-5, "who": r} if r == 0 else None, 0)
    assert obj == {"lr": 1e-5, "who": 0}
    m = torch.nn.Linear(4, 3)
    hvd.broadcast_parameters(m.state_dict(), 0)
    w = hvd.allgather(m.weight.detach().reshape(1, -1))
    assert torch.equal(w[0], w[-1])
    hvd.barrier()
    st = hvd.stats()
    assert st["fused_tensors"] >= 2, st
    # the hvd.callbacks.* that train.py does not use but Horovod ships
    from batchai_retinanet_horovod_coco_b200.callbacks import (LearningRateScheduleCallback, LearningRateWarmupCallback,
                                                                MetricAverageCallback)
    logs = {"loss": float(r), "mAP": 1.0 + r, "note": "text"}
    MetricAverageCallback().on_epoch_end(0, logs)
    assert abs(logs["loss"] - (n - 1) / 2) < 1e-9 and abs(logs["mAP"] - (1.0 + (n - 1) / 2)) < 1e-9 and logs["note"] == "text"

    class _T:
        lr = 0.8
    warm = LearningRateWarmupCallback(warmup_epochs=4)
    warm.trainer = _T()
    warm.on_epoch_begin(0)
    assert abs(_T.lr - 0.8) < 1e-12 and abs(warm.trainer.lr - 0.8 / n) < 1e-12     # starts at lr / size ...
    warm.on_epoch_begin(2)
    assert abs(warm.trainer.lr - 0.8 / n * (2 * (n - 1) / 4 + 1)) < 1e-12          # ... and ramps linearly to lr
    sched = LearningRateScheduleCallback(lambda e: 0.1 ** e, start_epoch=1)
    sched.trainer = _T()
    sched.on_epoch_begin(0)
    sched.on_epoch_begin(2)
    assert abs(sched.trainer.lr - 0.8 * 0.01) < 1e-12
    x = torch.tensor([1.0002, -3.0]) * (r + 1)
    y = hvd.allreduce(x, average=False, compression=hvd.Compression.fp16)
    want = sum(float(torch.tensor(1.0002 * (k + 1)).half()) for k in range(n))
    assert y.dtype == torch.float32 and abs(y[0].item() - want) < 1e-5, (y, want)   # values travel fp16-rounded
    hvd.timeline_stop()
elif mode == "ar_then_bcast":
    # An allreduce and a broadcast released by the SAME response: the root must not overwrite the shared
    # result area while slower ranks are still copying the allreduce result out of it (the broadcast stages
    # in the root's own region).  Rank 1 is slowed down by B200RT_FAULT=delay.
    from batchai_retinanet_horovod_coco_b200.parallel import runtime as _rt
    nel = 1 << 20
    for it in range(12):
        a = torch.full((nel,), float(r + 1 + it))
        b = torch.full((nel,), -7.0 - it) if r == 0 else torch.zeros(nel)
        ha = hvd.allreduce_async_(a, average=False, name=f"ar{it}")
        hb = _rt.enqueue(_rt.BROADCAST, f"bc{it}", b, b, nel, 0)
        ha.wait()
        hb.wait()
        want = float(sum(q + 1 + it for q in range(n)))
        assert bool((a == want).all()), (it, a.min().item(), a.max().item(), want)
        assert bool((b == -7.0 - it).all()), (it, b[:4])
elif mode == "mismatch":
    t = torch.zeros(4 if r == 0 else 5)
    try:
        hvd.allreduce_(t, name="bad")
    except RuntimeError as e:
        assert "mismatched shapes" in str(e), e
        print("MISMATCH_OK", r, flush=True)
    else:
        raise SystemExit(3)
elif mode == "stall":
    if r != 0:
        time.sleep(1.5)
    hvd.allreduce_(torch.ones(3), name="late")
    hvd.barrier()
    if r == 0:
        assert hvd.stats()["stall_warnings"] >= 1
elif mode == "die":
    if r == 1:
        os._exit(5)
    hvd.allreduce_(torch.ones(3), name="never")
print("WORKER_OK", r, flush=True)
hvd.shutdown()
