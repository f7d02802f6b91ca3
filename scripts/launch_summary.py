"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: last training step only."""
import csv, sys, collections
path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/launches.csv"
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
recs = list(csv.DictReader(lines))
idx = [i for i, x in enumerate(recs) if 'preprocess_kernel' in x['Kernel Name']][-1]
step = recs[idx:]
def ms(x):
    v = float(x['Metric Value'].replace(',', '')); u = x['Metric Unit']
    return v / 1e6 if u.startswith('ns') else v / 1e3 if u.startswith('us') else v
agg = collections.defaultdict(lambda: [0, 0.0])
for x in step:
    n = x['Kernel Name'].split('(')[0].replace('b200::', '')[:48]
    agg[n][0] += 1; agg[n][1] += ms(x)
tot = sum(v[1] for v in agg.values())
print(f"step: {len(step)} launches, {tot:.3f} ms")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print(f"{v[1]:8.3f} ms {100*v[1]/tot:5.1f}% {v[0]:4d}x  {k}")
if len(sys.argv) > 2:
    out = []
    for i, x in enumerate(step):
        n = x['Kernel Name'].split('(')[0].replace('b200::', '')
        n = n.replace('conv_igemm_kernel', 'conv').replace('wgrad_kernel', 'wgrad')[:12]
        out.append(f"{i}:{n}:{ms(x):.3f}")
    print(" ".join(out))
