"""Turn an .ncu-rep into a short markdown summary (key raw metrics + top stall lines)."""
import csv, subprocess, sys, io
rep, title = sys.argv[1], sys.argv[2]
kern = sys.argv[3] if len(sys.argv) > 3 else None     # optional kernel-name regex; the LAST matching launch is used
import re
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
ki = hdr.index("Kernel Name")
cands = [r for r in rows[2:] if kern is None or re.search(kern, r[ki])]
vals = cands[-1]
m = {h: (u, v) for h, u, v in zip(hdr, units, vals)}
keys = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
print(f"## {title}\n")
print(f"kernel: `{m.get('Kernel Name', ('', '?'))[1]}`\n")
print("| metric | value | unit |\n|---|---|---|")
for k in keys:
    for h in m:
        if h == k or h.endswith("." + k):
            print(f"| {k} | {m[h][1]} | {m[h][0]} |")
            break
st = [(h, v) for h, (u, v) in m.items() if "pcsamp_warps_issue_stalled" in h and "not_issued" not in h]
st = sorted(st, key=lambda x: -float(x[1].replace(",", "") or 0))[:6]
print("\nwarp stall samples (top): " + ", ".join(f"{h.split('stalled_')[-1]} {v}" for h, v in st))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
srows = list(csv.reader(io.StringIO(src)))
recs, si, ci, take = [], None, None, kern is None
for r in srows:                       # "Kernel Name" line, header, body per profiled launch; keep the last matching launch
    if len(r) >= 2 and r[0] == "Kernel Name":
        nm = r[1].replace("(bool)", "").replace("b200::", "")      # the source page spells template args differently
        take = kern is None or bool(re.search(kern, nm))
        si = None
        continue
    if not take:
        continue
    if "Source" in r and any(c.startswith("Warp Stall Sampling (All") for c in r):
        si, ci = r.index("Source"), [j for j, c in enumerate(r) if c.startswith("Warp Stall Sampling (All")][0]
        recs = []
        continue
    if si is not None and len(r) > max(si, ci):
        try:
            recs.append((int(r[ci] or 0), r[si].strip()))
        except ValueError:
            pass
tot = sum(s for s, _ in recs) or 1
print("\ntop SASS by stall samples:\n\n```")
for s, t in sorted(recs, key=lambda x: -x[0])[:10]:
    print(f"{100 * s / tot:5.1f}%  {t[:100]}")
print("```")
mn = {}
for _, t in recs:
    op = t.split()[0] if t and not t.startswith("@") else (t.split()[1] if len(t.split()) > 1 else "")
    for key in ("UTCHMMA", "UTMALDG", "LDTM", "UTCBAR", "SYNCS", "REDG", "RED.", "STG", "LDG"):
        if op.startswith(key):
            mn[key] = mn.get(key, 0) + 1
print("\nSASS mnemonics present: " + ", ".join(f"{k} x{v}" for k, v in sorted(mn.items())))
