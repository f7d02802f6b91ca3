"""Per-kernel SASS listings of libb200ops.so into profiles/sass/ (no GPU needed).  One file per kernel function with the
mnemonic counts that prove the Blackwell path (UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA load / store, LDTM =
tcgen05.ld, UTCBAR = tcgen05.commit, SYNCS = mbarrier, LDGMC / STG...MC = multimem) in a header comment."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "batchai_retinanet_horovod_coco_b200", "_native", "libb200ops.so")
out_dir = os.path.join(ROOT, "profiles", "sass")
os.makedirs(out_dir, exist_ok=True)
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
funcs, cur = collections.OrderedDict(), None
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        funcs[cur].append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line).rstrip())
keys = ("UTCHMMA", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "ELECT", "REDG", "RED.", "LDGMC",
        "STRONG.SYS", "HMMA", "ACQBULK", "PREEXIT", "UTMACMDFLUSH", "FENCE", "MULTICAST", "UCGABAR")
index = []
for mangled, lines in funcs.items():
    name = demangle(mangled)
    if re.search(r"_kernel<[3-7]>", name):        # world sizes 3..7 of the collectives: same code as <2> / <8>
        continue
    short = re.sub(r"\(.*", "", name).replace("void ", "").replace("b200::", "").replace("(anonymous namespace)::", "")
    fname = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_")[:80] + ".sass"
    cnt = collections.Counter()
    for l in lines:
        op = l.split("*/", 1)[-1].strip().lstrip("@!P0123456789UT ").split(" ")[0] if "*/" in l else ""
        for k in keys:
            if k in l.split("*/", 1)[-1][:60]:
                cnt[k] += 1
    with open(os.path.join(out_dir, fname), "w") as f:
        f.write(f"// {name}\n// {len(lines)} SASS instructions; " + ", ".join(f"{k} x{v}" for k, v in sorted(cnt.items())) + "\n")
        f.write("\n".join(lines) + "\n")
    index.append((short, len(lines), dict(cnt), fname))
with open(os.path.join(out_dir, "INDEX.md"), "w") as f:
    f.write("# SASS listings (cuobjdump -sass libb200ops.so, one file per kernel)\n\n| kernel | instructions | Blackwell mnemonics | file |\n|---|---|---|---|\n")
    for short, n, cnt, fname in index:
        f.write(f"| `{short}` | {n} | " + ", ".join(f"{k} x{v}" for k, v in sorted(cnt.items())) + f" | {fname} |\n")
print(len(index), "kernels")
