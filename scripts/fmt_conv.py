import sys, json
for l in sys.stdin:
    try:
        r = json.loads(l)
    except Exception:
        print(l.strip()); continue
    print(r["name"], "fprop %.3f (%.0f TF) dgrad %.3f wgrad %.3f | cudnn %.3f %.3f %.3f" % (
        r["fprop_ms"], r["fprop_tflops"], r["dgrad_ms"], r["wgrad_ms"], r["cudnn_fprop_ms"], r["cudnn_dgrad_ms"], r["cudnn_wgrad_ms"]))
