import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
from batchai_retinanet_horovod_coco_b200.parallel.p2p import P2PComm

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
comm = P2PComm.from_torch_distributed(dev)
N = 6_000_000
bucket = comm.symmetric_bucket(N)
cases = [(1, 1000), (1, 100_000), (2, 100_000), (2, N), (0, 4096), (0, N), (3, 100_000), (3, N)]
if rank == 0:
    print("P2P_REGION", "vmm+multicast" if comm.mc_base else "cuda-ipc")
if comm.mc_base:      # NVLS: in-switch reduce + multicast store (algo 4), and the auto choice with a shared result buffer
    cases += [(4, 1000), (4, 100_000), (4, N), (-1, N)]
for algo, n in cases:
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    x = torch.randn(N, device=dev, generator=g)
    bucket.copy_(x)
    pre = 0.5 + rank
    comm.prescale_local.fill_(pre)
    out = comm.result if algo in (3, 4, -1) else torch.zeros(N, device=dev)
    out.zero_()
    torch.cuda.synchronize(); dist.barrier()
    comm.allreduce_avg(bucket, out, n, prescale=True, algo=max(algo, 0))
    torch.cuda.synchronize()
    ref = x * pre
    dist.all_reduce(ref)
    ref /= world
    n4 = (n + 3) // 4 * 4
    err = (out[:n4] - ref[:n4]).abs().max().item()
    assert err < 1e-4 * (1 + ref.abs().max().item()), (rank, algo, n, err)
    assert n4 == N or out[n4:].abs().max().item() == 0
# staged allreduce (the engine's overlapped path): raw slices scattered from a side stream in pieces, then the weighted
# local reduce + gather; fp32 wire must match the plain allreduce to rounding, bf16 wire to bf16 rounding of the peers
side = torch.cuda.Stream(device=dev)
for wire, n in (("fp32", 100_000), ("fp32", N), ("bf16", N), ("fp32", 4096)):
    os.environ["B200_AR_WIRE"] = wire
    g = torch.Generator(device=dev).manual_seed(300 + rank)
    x = torch.randn(N, device=dev, generator=g)
    bucket.copy_(x)
    pre = 0.25 + 0.5 * rank
    comm.prescale_local.fill_(pre)
    comm.result.zero_()
    torch.cuda.synchronize(); dist.barrier()
    n4 = (n + 3) // 4 * 4
    cuts = [0, (n4 // 3) // 4 * 4, (2 * n4 // 3) // 4 * 4, n4]
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        comm.scatter(n, lo, hi, stream=side)
    torch.cuda.current_stream(dev).wait_stream(side)
    comm.allreduce_staged_avg(n, prescale=True)
    torch.cuda.synchronize()
    ref = x * pre
    dist.all_reduce(ref)
    ref /= world
    err = (comm.result[:n4] - ref[:n4]).abs().max().item()
    tol = (1e-4 if wire == "fp32" else 2e-2) * (1 + ref.abs().max().item())
    assert err < tol, (rank, wire, n, err)
    chk = comm.result[:n4].clone()
    dist.broadcast(chk, 0)
    assert torch.equal(chk, comm.result[:n4]), "staged allreduce: ranks hold different results"
os.environ["B200_AR_WIRE"] = "fp32"
t = torch.full((1_000_003,), float(rank + 1), device=dev)
comm.broadcast_(t, 0)
assert (t == 1).all()
t = torch.full((12345,), float(rank + 1), device=dev)
comm.allreduce_sum_(t)
assert abs(t[0].item() - world * (world + 1) / 2) < 1e-5
# flag-protocol stress: ranks drift apart by random host delays, collectives are issued back to back
# without host syncs and the bucket is rewritten right after each one -- a rank that left a barrier early, or an epoch
# that was reused, shows up as a wrong average
import random, time
rnd = random.Random(1000 + rank)
err = torch.zeros((), device=dev)
for it in range(300):
    if rnd.random() < 0.25:
        time.sleep(rnd.random() * 0.003)
    n = (1024, 4096, 262144, 1 << 20)[it % 4]
    bucket[:n].fill_(float(rank + it))
    comm.allreduce_avg(bucket, comm.result, n, prescale=False, algo=(0, 1, 2, 3)[(it // 4) % 4])
    err = torch.maximum(err, (comm.result[:n] - (it + (world - 1) / 2)).abs().max())
torch.cuda.synchronize()
assert err.item() < 1e-4, (rank, err.item())
# the same stress for the staged path: scatter and reduce back to back, staging and bucket rewritten immediately
err = torch.zeros((), device=dev)
for it in range(200):
    if rnd.random() < 0.25:
        time.sleep(rnd.random() * 0.003)
    n = (4096, 262144, 1 << 20)[it % 3]
    bucket[:n].fill_(float(rank + it))
    side.wait_stream(torch.cuda.current_stream(dev))
    comm.scatter(n, 0, n, stream=side)
    torch.cuda.current_stream(dev).wait_stream(side)
    comm.allreduce_staged_avg(n, prescale=False)
    err = torch.maximum(err, (comm.result[:n] - (it + (world - 1) / 2)).abs().max())
torch.cuda.synchronize()
assert err.item() < 1e-4, (rank, "staged", err.item())
# exposed time of the staged path's foreground half (reduce + gather) next to the full allreduce
for nbytes in (1 << 20, 1 << 24, N * 4):
    n = nbytes // 4
    res = {}
    for name in ("full", "staged"):
        def call():
            if name == "full":
                comm.allreduce_avg(bucket, comm.result, n, prescale=False)
            else:
                comm.allreduce_staged_avg(n, prescale=False)
        comm.scatter(n, 0, n)
        for _ in range(3):
            call()
        torch.cuda.synchronize(); dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(10):
            call()
        e.record(); torch.cuda.synchronize()
        ms = torch.tensor([s.elapsed_time(e) / 10], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        res[name] = ms.item()
    if rank == 0:
        print(f"P2P_STAGED bytes={nbytes} world={world} full_ms={res['full']:.4f} staged_foreground_ms={res['staged']:.4f}")
# bandwidth (device timed, max over ranks)
for nbytes in (1 << 12, 1 << 16, 1 << 20, 1 << 24, N * 4):
    n = nbytes // 4
    out = comm.result          # shared result buffer: lets the auto choice use push / NVLS
    for _ in range(3):
        comm.allreduce_avg(bucket, out, n, prescale=False)
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10):
        comm.allreduce_avg(bucket, out, n, prescale=False)
    e.record(); torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e) / 10], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"P2P_BW bytes={nbytes} world={world} ms={ms.item():.4f} algbw_GBs={nbytes / ms.item() / 1e6:.1f} "
              f"busbw_GBs={nbytes / ms.item() / 1e6 * 2 * (world - 1) / world:.1f}")
comm.close()
if rank == 0:
    print("P2P_OK")
dist.destroy_process_group()
