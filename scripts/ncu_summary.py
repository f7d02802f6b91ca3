"""Turn an .ncu-rep into a short markdown summary (key raw metrics + top stall lines)."""
import csv, subprocess, sys, io
rep, title = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
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
srows = list(csv.reader(io.StringIO(src)))[2:]
recs = [(int(r[2] or 0), r[1].strip()) for r in srows if len(r) > 5]
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
