#!/usr/bin/env python
"""Headline benchmark: RetinaNet-R50 training images/sec on N B200s (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl b200|library|reference]

* ``--impl b200`` (default): this repo's native engine -- hand-written sm_100a kernels
  (tcgen05 implicit-GEMM convs, fused losses/optimizer, P2P NVLink allreduce).
* ``--impl library``: the same model in PyTorch eager with cuDNN + NCCL/DDP (BASELINE.md section 5
  stand-in baseline; not the product).
* ``--impl reference``: the unmodified reference from ``baseline/_ref``.  It cannot be installed
  offline (no setup.py, needs TF1/Keras2.1/Horovod/MPI), so this prints
  ``{"impl": "reference", "unavailable": ...}`` and exits 0.

Config = BASELINE.json config 2: ResNet-50 FPN, 800x1333 synthetic COCO-shaped batches, bf16,
data parallel, weak scaling (fixed per-GPU batch).  Synthetic data, random-init weights.
Timed with CUDA events on the device after warm-up, max over ranks; ``e2e`` additionally copies
every step's inputs from pinned host memory and reads the loss back.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "library", "reference"])
    ap.add_argument("--backbone", default="resnet50")
    ap.add_argument("--batch", type=int, default=16, help="per-GPU batch (weak scaling)")
    ap.add_argument("--height", type=int, default=800)
    ap.add_argument("--width", type=int, default=1333)
    ap.add_argument("--classes", type=int, default=80)
    ap.add_argument("--boxes", type=int, default=12, help="synthetic GT boxes per image")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--with-library", action="store_true",
                    help="also time the library stand-in (PyTorch + cuDNN, CUDA-graphed on 1 GPU, DDP/NCCL on N) in the "
                         "same process on the same box and report value / library as vs_library")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.proc = None
        self.gpu = gpu_index
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(gpu_index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def synth_batches(args, n_sets: int, seed: int):
    """COCO-shaped synthetic batches in pinned host memory: uint8 BGR images + (G,5) boxes."""
    import torch
    g = torch.Generator().manual_seed(seed)
    sets = []
    B, H, W, G = args.batch, args.height, args.width, args.boxes
    for _ in range(n_sets):
        img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g)
        xy = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.7, H * 0.7])
        wh = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.3, H * 0.3]) + 24.0
        lab = torch.randint(0, args.classes, (B, G, 1), generator=g).float()
        gt = torch.cat([xy, xy + wh, lab], dim=2)
        cnt = torch.full((B,), G, dtype=torch.int32)
        hw = torch.tensor([[H, W]] * B, dtype=torch.int32)
        tensors = [img, gt, cnt, hw]
        if torch.cuda.is_available():
            tensors = [t.pin_memory() for t in tensors]
        sets.append(tuple(tensors))
    return sets


def main():
    args = parse()
    if args.impl == "reference":
        print(json.dumps({"impl": "reference",
                          "unavailable": "reference has no setup.py/pyproject and needs TensorFlow1/Keras2.1/"
                                         "Horovod/MPI/h5py/pycocotools, none installable offline (see DESIGN.md)"}))
        return 0

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.backends.cudnn.benchmark = True

    if args.impl == "library":
        from batchai_retinanet_horovod_coco_b200.engine.library import LibraryTrainer
        trainer = LibraryTrainer(args.backbone, args.classes, device=dev, distributed=world > 1)
    else:
        from batchai_retinanet_horovod_coco_b200 import models
        from batchai_retinanet_horovod_coco_b200.engine import native_engine_class
        model = models.backbone(args.backbone).retinanet(args.classes)
        trainer = native_engine_class(model)(model=model, device=dev, batch=args.batch,
                                            image_hw=(args.height, args.width), distributed=world > 1)

    from batchai_retinanet_horovod_coco_b200.ops import native
    host_sets = synth_batches(args, n_sets=4, seed=1234 + rank)
    dev_sets = [tuple(t.to(dev) for t in s) for s in host_sets]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step_async = getattr(trainer, "train_step_async", None)

    def run(n, sets, read_loss):
        last = None
        for i in range(n):
            if read_loss and step_async is not None:
                # public API: the handle's loss_cpu() is the device->host read of the step's result (loss triple);
                # it is ready after the forward half, so the next step's H2D copy overlaps this step's backward
                last = step_async(*sets[i % len(sets)]).loss_cpu()
                continue
            last = trainer.train_step(*sets[i % len(sets)])
            if read_loss:
                last = last.cpu()  # device->host read of the step's result (loss triple)
        return last

    def timed(n, sets, read_loss):
        barrier()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        out = run(n, sets, read_loss)
        end.record()
        barrier()
        ms = torch.tensor([start.elapsed_time(end)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out

    # warm-up (untimed), then the device-timed region with clocks sampled during it
    run(args.warmup, dev_sets, False)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    native_ok = os.path.exists(native.lib_path())
    l0 = native.launch_count() if native_ok else 0
    ms, last = timed(args.steps, dev_sets, False)
    launches = (native.launch_count() - l0) if native_ok else 0
    clocks = sampler.stop() if sampler else None
    loss_val = [float(v) for v in last.float().cpu()]

    e2e = None
    if not args.no_e2e:
        run(2, host_sets, True)
        ms_e, _ = timed(args.steps, host_sets, True)
        h2d = sum(t.numel() * t.element_size() for t in host_sets[0])
        e2e = {"value": args.batch * world * args.steps / (ms_e / 1e3), "unit": "images/sec",
               "ms_per_step": ms_e / args.steps, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 12}

    # Every rank must hold bit-identical master weights after the same averaged updates (Horovod's contract,
    # reference train.py:103,111): compare a checksum of the flat parameter buffer across ranks.
    replicas_equal = None
    P = getattr(trainer, "P", None)
    if P is not None:
        n_tr = getattr(trainer, "n_train", P.numel())
        bits = P[:n_tr].view(torch.int32).to(torch.int64)
        chk = torch.stack([bits.sum(), (bits * (torch.arange(n_tr, device=dev) % 8191 + 1)).sum()])   # int64, exact
        if world > 1:
            lo, hi = chk.clone(), chk.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            replicas_equal = bool((lo == hi).all().item())
        else:
            replicas_equal = True

    library = None
    if args.with_library and args.impl == "b200":
        from batchai_retinanet_horovod_coco_b200.engine.library import LibraryTrainer
        del trainer
        torch.cuda.empty_cache()
        trainer = LibraryTrainer(args.backbone, args.classes, device=dev, distributed=world > 1)
        step_async = None
        lsteps = max(5, args.steps // 4)
        run(6, dev_sets, False)
        ms_l, _ = timed(lsteps, dev_sets, False)
        library = {"value": args.batch * world * lsteps / (ms_l / 1e3), "ms_per_step": ms_l / lsteps, "steps": lsteps,
                   "what": "same model in PyTorch (cuDNN bf16 autocast, channels-last, fused capturable Adam, whole-step "
                           "CUDA graph on 1 GPU / DDP+NCCL on N); same box, same process"}

    if rank == 0:
        value = args.batch * world * args.steps / (ms / 1e3)
        line = {
            "metric": "RetinaNet-R50 train imgs/sec" if args.backbone == "resnet50" else f"RetinaNet-{args.backbone} train imgs/sec",
            "impl": args.impl, "value": value, "unit": "images/sec", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic (uint8 COCO-shaped images + boxes, random-init weights)",
            "config": {"model": f"retinanet-{args.backbone}-fpn", "global_batch": args.batch * world,
                       "per_gpu_batch": args.batch, "image": [args.height, args.width], "seq_len": None,
                       "parallelism": f"dp{world}", "optimizer": "adam lr1e-5 clipnorm1e-3",
                       "l2_policy": "inputs+activations per step (>2 GB) exceed the 126 MB L2; 4 rotating input sets"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "loss": loss_val,
            "replicas_equal": replicas_equal,
            "library": library, "vs_library": (value / library["value"]) if library else None,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
