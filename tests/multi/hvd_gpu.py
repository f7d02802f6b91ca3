"""Horovod-style API on CUDA tensors over 2 ranks: DistributedOptimizer around a plain torch.optim.SGD averages the
gradients through the fused P2P path while backward is still running (hooks -> side stream), Compression.fp16 halves the
wire, handles are stream-ordered."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import batchai_retinanet_horovod_coco_b200.parallel as hvd

hvd.init()
rank, world = hvd.rank(), hvd.size()
torch.cuda.set_device(hvd.local_rank())
dev = torch.device("cuda", hvd.local_rank())
torch.manual_seed(0)
model = torch.nn.Sequential(torch.nn.Linear(512, 1024), torch.nn.ReLU(), torch.nn.Linear(1024, 1024), torch.nn.ReLU(),
                            torch.nn.Linear(1024, 10)).to(dev)
hvd.broadcast_parameters(model.state_dict(), 0)
opt = hvd.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1), named_parameters=model.named_parameters())
g = torch.Generator().manual_seed(10 + rank)
x = torch.randn(64, 512, generator=g).to(dev)
y = torch.randint(0, 10, (64,), generator=g).to(dev)
loss = torch.nn.functional.cross_entropy(model(x), y)
loss.backward()                      # hooks enqueue every gradient as autograd produces it
local = [p.grad.clone() for p in model.parameters()]     # NOTE: taken before synchronize(): may already be averaged
opt.synchronize()
torch.cuda.synchronize()
# reference: average of the per-rank gradients recomputed without hooks
ref_model = torch.nn.Sequential(torch.nn.Linear(512, 1024), torch.nn.ReLU(), torch.nn.Linear(1024, 1024), torch.nn.ReLU(),
                                torch.nn.Linear(1024, 10)).to(dev)
ref_model.load_state_dict(model.state_dict())
torch.nn.functional.cross_entropy(ref_model(x), y).backward()
for p, q in zip(model.parameters(), ref_model.parameters()):
    want = hvd.allreduce(q.grad, average=True)
    assert torch.allclose(p.grad, want, rtol=1e-5, atol=1e-7), (p.grad - want).abs().max()
f = hvd._state["gpu_fusion"]
assert f is not None and f.done_prev is not None, "the fused GPU path was not used"
opt.step()
chk = hvd.allgather(torch.cat([p.detach().reshape(-1)[:8] for p in model.parameters()]).reshape(1, -1).cpu())
assert torch.equal(chk[0], chk[-1]), "replicas diverged"
# stream-ordered handle + poll
t = torch.full((1 << 20,), float(rank + 1), device=dev)
h = hvd.allreduce_async_(t, average=False, name="big")
h.wait()
assert abs(t[123].item() - world * (world + 1) / 2) < 1e-5 and h.poll()
# Compression.fp16: result equals the fp32 one to bf16 rounding of the PEERS' contributions
v = torch.randn(1 << 18, device=dev, generator=torch.Generator(device=dev).manual_seed(5 + rank))
exact = hvd.allreduce(v, average=True)
half = hvd.allreduce(v, average=True, compression=hvd.Compression.fp16)
err = (half - exact).abs().max().item()
assert 0 < err < 2e-2 * exact.abs().max().item(), err
hvd.barrier()
print("HVD_GPU_OK", rank, flush=True)
hvd.shutdown()
