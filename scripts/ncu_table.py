"""One markdown row per kernel of an .ncu-rep (last launch of each name): time, DRAM bytes and %, tensor %, registers,
shared memory, top stall reason.  Usage: python scripts/ncu_table.py rep.ncu-rep > profiles/kernels_vN.md"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
def col(name):
    for i, h in enumerate(hdr):
        if h == name or h.endswith("." + name):
            return i
    return None
ki = hdr.index("Kernel Name")
ALL = len(sys.argv) > 2 and sys.argv[2] == "all"          # every launch in order instead of the last one per name
last = {}
for n, r in enumerate(rows[2:]):
    last[(n, r[ki]) if ALL else r[ki]] = r
stall_cols = [(i, h.split("stalled_")[-1]) for i, h in enumerate(hdr) if "pcsamp_warps_issue_stalled" in h and "not_issued" not in h]
_SCALE = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6,
          "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
def f(r, name, scale=1.0, fmt="{:.1f}"):
    c = col(name)
    try:
        # ncu picks the unit per column (us or ms, Mbyte or Gbyte): normalise times to us and sizes to MB
        return fmt.format(float(r[c].replace(",", "")) * scale * _SCALE.get(units[c], 1.0))
    except Exception:
        return "-"
print("| kernel | us | DRAM rd MB | DRAM wr MB | DRAM % | L2 % | tensor % | regs | smem KB | top stalls |")
print("|---|---|---|---|---|---|---|---|---|---|")
order = last.items() if ALL else sorted(last.items(), key=lambda kv: -float(kv[1][col("gpu__time_duration.sum")].replace(",", "") or 0))
for name, r in order:
    name = f"{name[0]}: {name[1]}" if ALL else name
    st = sorted(((float(r[i].replace(",", "") or 0), n) for i, n in stall_cols), reverse=True)[:2]
    tot = sum(float(r[i].replace(",", "") or 0) for i, _ in stall_cols) or 1
    short = name.replace("b200::", "").replace("void ", "").split("(")[0]
    print(f"| `{short}` | {f(r, 'gpu__time_duration.sum')} | {f(r, 'dram__bytes_read.sum')} | {f(r, 'dram__bytes_write.sum')} | "
          f"{f(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')} | {f(r, 'lts__throughput.avg.pct_of_peak_sustained_elapsed')} | "
          f"{f(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active')} | {f(r, 'launch__registers_per_thread', fmt='{:.0f}')} | "
          f"{f(r, 'launch__shared_mem_per_block_dynamic')} | " + ", ".join(f"{n} {100 * v / tot:.0f}%" for v, n in st) + " |")
