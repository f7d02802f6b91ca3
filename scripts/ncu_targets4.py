"""The nine conv launches of one res4 identity block (8x50x84: 263 pixel tiles on 148 SMs, the layers furthest from both
roofs in profiles/layer_times_b8_v5.md) between cudaProfilerStart/Stop for one `ncu --set full` capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer
from bench_configs_common import batch, dev

B, H, W = 8, 800, 1333
nt = NativeTrainer("resnet50", 80, device=dev, batch=B, image_hw=(H, W), use_graph=False)
b = batch(B, H, W)
for _ in range(2):
    nt.train_step(*b)
p = nt._plan(B, H, W)
a, d, gb = p["acts"][2][1], nt.blocks[2][1], p["gb"][2]
torch.cuda.synchronize()
torch.cuda.profiler.start()
nt._fprop(d["a"], a["x"], a["a"], relu=True)                    # 0: 1x1 1024->256
nt._fprop(d["b"], a["a"], a["b"], relu=True)                    # 1: 3x3 256->256
nt._fprop(d["c"], a["b"], a["o"], relu=True, add=a["x"])        # 2: 1x1 256->1024 + residual
nt._dgrad(d["c"], gb["g"], gb["gb"], mask=a["b"])               # 3: dgrad of 256->1024 (K = 1024)
nt._dgrad(d["b"], gb["gb"], gb["ga"], mask=a["a"])              # 4: dgrad 3x3
nt._dgrad(d["a"], gb["ga"], gb["g"], mask=a["x"], add=gb["g"])  # 5: dgrad of 1024->256 (N = 1024) + residual + mask
nt._wgrad(d["a"], [a["x"]], [gb["ga"]])                         # 6
nt._wgrad(d["b"], [a["a"]], [gb["gb"]])                         # 7
nt._wgrad(d["c"], [a["b"]], [gb["g"]])                          # 8
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")
