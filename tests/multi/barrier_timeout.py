"""Fault injection for the in-kernel flag barrier: rank 1 skips a collective, rank 0's kernel must give up after
B200_AR_TIMEOUT_S instead of spinning for ever, and P2PComm.check() must name the missing rank."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ["B200_AR_TIMEOUT_S"] = "1.0"
import torch, torch.distributed as dist
from batchai_retinanet_horovod_coco_b200.parallel.p2p import P2PComm

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
comm = P2PComm.from_torch_distributed(dev)
bucket = comm.symmetric_bucket(1 << 20)
bucket.fill_(float(rank + 1))
comm.allreduce_avg(bucket, comm.result, 1 << 16, prescale=False)       # a healthy collective first
torch.cuda.synchronize()
comm.check()
assert abs(comm.result[0].item() - (world + 1) / 2) < 1e-6
dist.barrier()
if rank == 0:
    t0 = time.time()
    comm.allreduce_avg(bucket, comm.result, 1 << 16, prescale=False)   # rank 1 never joins this one
    torch.cuda.synchronize()
    dt = time.time() - t0
    assert dt < 20, f"barrier did not time out ({dt:.1f} s)"
    try:
        comm.check()
    except RuntimeError as e:
        assert "rank 1" in str(e), e
        print("TIMEOUT_OK", f"{dt:.2f}s", flush=True)
    else:
        raise SystemExit("check() did not report the timed-out barrier")
else:
    time.sleep(6)
sys.stdout.flush()
os._exit(0)        # the communicator is broken on purpose: no orderly close
