"""Per-launch device time of every conv / wgrad call of one eager training step (events around each call), with the
TFLOP/s and the minimum-traffic GB/s it achieved.  Shows which layers sit far from both roofs."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench_configs_common import batch, dev
from batchai_retinanet_horovod_coco_b200.ops import native
from batchai_retinanet_horovod_coco_b200.engine import native_engine_class
from batchai_retinanet_horovod_coco_b200.models.retinanet import RetinaNet

B = int(os.environ.get("B", 8))
H, W = 800, 1333
model = RetinaNet(80, "resnet50")
tr = native_engine_class(model)(model=model, device=dev, batch=B, image_hw=(H, W), use_graph=False)
b = batch(B, H, W)
rec, on = [], [False]
_conv, _wgrad = native.conv_raw, native.wgrad_raw


def _timed(fn, meta):
    if not on[0]:
        return fn()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); fn(); e.record()
    rec.append((meta, s, e))


def conv_raw(mode, problems, weight, K, ldi, Nrows, ldo, KH, KW, stride, pad_h, pad_w, **kw):
    px = sum(p[1].shape[0] * p[1].shape[1] * p[1].shape[2] for p in problems)
    taps = KH * KW if not (mode == 1 and stride == 2) else max(1, KH * KW // 4)
    by = sum(p[0].numel() * 2 + p[1].numel() * (4 if kw.get("out_fp32") else 2) + sum(t.numel() * 2 for t in p[2:4] if t is not None) for p in problems)
    meta = dict(kind="dgrad" if mode else "fprop", shape=f"{len(problems)}x[{problems[0][0].shape[1]}x{problems[0][0].shape[2]}] {K}->{Nrows} {KH}x{KW}/{stride}",
                gflop=2e-9 * px * K * Nrows * taps, mb=(by + weight.numel() * 2) / 1e6)
    return _timed(lambda: _conv(mode, problems, weight, K, ldi, Nrows, ldo, KH, KW, stride, pad_h, pad_w, **kw), meta)


def wgrad_raw(problems, dw, Cin, ldx, Cout, ldy, KH, KW, stride, pad_h, pad_w, **kw):
    px = sum(p[1].shape[0] * p[1].shape[1] * p[1].shape[2] for p in problems)
    by = sum(p[0].numel() * 2 + p[1].numel() * 2 for p in problems) + dw.numel() * 4
    meta = dict(kind="wgrad", shape=f"{len(problems)}x[{problems[0][0].shape[1]}x{problems[0][0].shape[2]}] {Cin}->{Cout} {KH}x{KW}/{stride}",
                gflop=2e-9 * px * Cin * Cout * KH * KW, mb=by / 1e6)
    return _timed(lambda: _wgrad(problems, dw, Cin, ldx, Cout, ldy, KH, KW, stride, pad_h, pad_w, **kw), meta)


native.conv_raw, native.wgrad_raw = conv_raw, wgrad_raw
for _ in range(3):
    tr.train_step(*b)
on[0] = True
tr.train_step(*b)
torch.cuda.synchronize()
rows = []
for meta, s, e in rec:
    ms = s.elapsed_time(e)
    rows.append(dict(meta, ms=round(ms, 4), tflops=round(meta["gflop"] / ms, 1), gbs=round(meta["mb"] / ms, 0)))
tot = sum(r["ms"] for r in rows)
agg = {}
for r in rows:
    k = (r["kind"], r["shape"])
    a = agg.setdefault(k, dict(kind=r["kind"], shape=r["shape"], n=0, ms=0.0, gflop=0.0, mb=0.0))
    a["n"] += 1; a["ms"] += r["ms"]; a["gflop"] += r["gflop"]; a["mb"] += r["mb"]
print(f"batch {B}: {len(rows)} conv/wgrad launches, {tot:.2f} ms (event-timed one by one, eager)")
print("| kind | problems x [HxW] K->N taps/stride | n | ms | % | TFLOP/s | GB/s (min traffic) |\n|---|---|---|---|---|---|---|")
for a in sorted(agg.values(), key=lambda a: -a["ms"]):
    print(f"| {a['kind']} | {a['shape']} | {a['n']} | {a['ms']:.3f} | {100 * a['ms'] / tot:.1f} | {a['gflop'] / a['ms']:.0f} | {a['mb'] / a['ms']:.0f} |")
