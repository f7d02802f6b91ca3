"""Gradient-allreduce bandwidth sweep (BASELINE.json config 5): the fused P2P kernel vs NCCL on the
same buffers, 1 KB .. 1 GB, device-timed with CUDA events, max over ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_allreduce.py

Prints one JSON line per size on rank 0: algorithm bandwidth (bytes / time), bus bandwidth
(x 2(N-1)/N) and the fraction of the NVLink roofline (770 GB/s measured peer copy per direction:
the fused reduce+push algorithm moves (N-1)/N * S bytes per direction per GPU)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from batchai_retinanet_horovod_coco_b200.parallel.p2p import P2PComm

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
comm = P2PComm.from_torch_distributed(dev)
MAXB = int(os.environ.get("MAX_BYTES", 1 << 30))
N = MAXB // 4
bucket = comm.symmetric_bucket(N)
bucket.normal_()
out = torch.empty(N, device=dev)
nccl_buf = torch.randn(N, device=dev)
LINK = 770e9


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record(); torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e) / iters], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item()


size = 1 << 10
while size <= MAXB:
    n = size // 4
    iters = 50 if size <= (1 << 22) else (10 if size <= (1 << 27) else 4)
    row = {"bytes": size, "world": world}
    for name, algo in (("oneshot", 1), ("twoshot", 2), ("push", 3)):
        if algo == 1 and size > (8 << 20):
            continue
        dst = comm.result if algo == 3 else out
        ms = timeit(lambda: comm.allreduce_avg(bucket, dst, n, prescale=True, algo=algo), iters)
        row[name + "_us"] = ms * 1e3
    if comm.mc_base:
        # NVLS: in-switch reduce + multicast store; `nvls_scaled` also pays the local in-place clip-factor pass
        row["nvls_us"] = timeit(lambda: comm.allreduce_avg(bucket, comm.result, n, prescale=True, algo=4), iters) * 1e3
        comm.prescale_local.fill_(0.999)
        row["nvls_scaled_us"] = timeit(lambda: comm.allreduce_avg(bucket, comm.result, n, prescale=True, algo=4), iters) * 1e3
        comm.prescale_local.fill_(1.0)
        bucket.normal_()
    best = min(v for k, v in row.items() if k.endswith("_us"))
    # staged path (the engine's default): the scatter runs beside backward; what stays on the critical path is the
    # flag barrier + local weighted reduce + all-gather through the switch.  Receives (N-1)/N * S per GPU.
    comm.scatter(n, 0, n)
    row["staged_scatter_us_hidden"] = timeit(lambda: comm.scatter(n, 0, n), iters) * 1e3
    fg = timeit(lambda: comm.allreduce_staged_avg(n, prescale=True), iters) * 1e3
    row["staged_foreground_us"] = fg
    row["staged_foreground_frac_of_nvlink_roofline"] = (world - 1) / world * size / LINK * 1e6 / fg
    ms_n = timeit(lambda: dist.all_reduce(nccl_buf[:n]), iters)
    row["nccl_us"] = ms_n * 1e3
    row["p2p_algbw_GBs"] = size / best / 1e3
    row["p2p_busbw_GBs"] = row["p2p_algbw_GBs"] * 2 * (world - 1) / world
    row["nccl_busbw_GBs"] = size / (ms_n * 1e3) / 1e3 * 2 * (world - 1) / world
    # per direction per GPU: unicast reduce + push moves (N-1)/N * S in and out; with NVLS the floor is S / LINK
    roof_us = (world - 1) / world * size / LINK * 1e6
    row["frac_of_nvlink_roofline"] = roof_us / best
    row["speedup_vs_nccl"] = ms_n * 1e3 / best
    if rank == 0:
        print(json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in row.items()}), flush=True)
    size *= 4
comm.close()
dist.destroy_process_group()
