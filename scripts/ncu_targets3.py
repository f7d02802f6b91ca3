"""Every kernel family of the step and of the prediction graph, a few launches each, between cudaProfilerStart/Stop for
ONE `ncu --set full --profile-from-start off` capture (round 2 evidence: conv variants incl. halo / kc modes, wgrad,
stem, pool, upsample, losses, target assignment, optimizer, weight refresh, detection tail)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from batchai_retinanet_horovod_coco_b200.engine.trainer import NativeTrainer, CAFFE_MEAN_BGR
from batchai_retinanet_horovod_coco_b200.ops import native

dev, bf = torch.device("cuda", 0), torch.bfloat16
B, H, W, G = 8, 800, 1333, 12
nt = NativeTrainer("resnet50", 80, device=dev, batch=B, image_hw=(H, W), use_graph=False)
g = torch.Generator().manual_seed(3)
img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g).to(dev)
xy = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.7, H * 0.7])
wh = torch.rand(B, G, 2, generator=g) * torch.tensor([W * 0.3, H * 0.3]) + 24.0
gt = torch.cat([xy, xy + wh, torch.randint(0, 80, (B, G, 1), generator=g).float()], dim=2).to(dev)
cnt = torch.full((B,), G, dtype=torch.int32, device=dev)
hw = torch.tensor([[H, W]] * B, dtype=torch.int32, device=dev)
for _ in range(2):
    nt.train_step(img, gt, cnt, hw)
p = nt._plan(B, H, W)
# make the classification logits non-degenerate so that the detection tail has candidates to sort and suppress
for l in range(5):
    p["cls_o"][l].copy_((torch.randn_like(p["cls_o"][l].float()) * 2.0 - 2.5).to(bf))
torch.cuda.synchronize()
torch.cuda.profiler.start()
native.preprocess(img, p["img"], CAFFE_MEAN_BGR, hw)
native.stem_conv(p["img"], nt.stem_wp, nt.stem.shift, p["stem_out"])
native.bn_relu_maxpool_fwd(p["stem_out"], nt.ones64, nt.zeros64, p["pool"], p["pool_idx"], *p["pool_pad"])
a0 = p["acts"][0][1]          # res2b: identity block at 200x334
d0 = nt.blocks[0][1]
nt._fprop(d0["a"], a0["x"], a0["a"], relu=True)                 # 1x1 256->64   (kc boxes, resident weights)
nt._fprop(d0["b"], a0["a"], a0["b"], relu=True)                 # 3x3 64->64    (halo boxes, resident weights)
nt._fprop(d0["c"], a0["b"], a0["o"], relu=True, add=a0["x"])    # 1x1 64->256 + residual
gb = p["gb"][0]
nt._dgrad(d0["c"], gb["g"], gb["gb"], mask=a0["b"])
nt._dgrad(d0["a"], gb["ga"], gb["g"], mask=a0["x"], add=gb["g"])  # residual + mask
nt._wgrad(d0["b"], [a0["a"]], [gb["gb"]])
native.upsample_add(p["fpn"]["m4"], p["fpn"]["c5r"], p["fpn"]["m4"])
nt._fprop_group(nt.cls_tower[0], p["fpn"]["P"], p["cls_t"][0], relu=True)     # head fprop, 5 levels
nt._losses(p, gt, cnt, hw)
nt._wgrad(nt.cls_out, p["cls_t"][3], p["dcls_o"])
nt._dgrad_group(nt.cls_out, p["dcls_o"], p["dt"][0], masks=p["cls_t"][3], colsums=[nt._bias_grad(nt.cls_tower[3])] * 5)
nt._wgrad(nt.cls_tower[3], p["cls_t"][2], p["dt"][0], dbias=False)
native.upsample_bwd_add(p["dfpn"]["m3"], p["dfpn"]["m4"])
native.bn_relu_maxpool_bwd(p["pool_idx"], nt.ones64, gb["gx"], p["dstem"], *p["pool_pad"])
native.stem_wgrad(p["img"], p["dstem"], nt.stem.scale, nt.stem.wview(nt.G))
nt._optimizer_step()
out = native.filter_detections_levels(p["cls_o"], p["reg_o"], 80, p["anchors"], (H, W), 0.05, 0.5, 300)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done", [tuple(o.shape) for o in out])
