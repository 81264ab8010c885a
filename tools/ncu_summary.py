"""Markdown summary of an .ncu-rep: key metrics per kernel + the most-sampled instructions (needs -lineinfo builds).
usage: python tools/ncu_summary.py report.ncu-rep > profiles/xyz.md"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
METRICS = [
    ("gpu__time_duration.sum", "duration"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thr"), ("launch__shared_mem_per_block_dynamic", "dyn smem"),
    ("sm__cycles_elapsed.max", "cycles"), ("smsp__cycles_active.avg", "active cyc/SMSP"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM thr %"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM thr %"),
    ("dram__bytes_read.sum", "DRAM rd"), ("dram__bytes_write.sum", "DRAM wr"), ("lts__t_bytes.sum", "L2 bytes"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
]
print(f"# ncu summary of `{rep.split('/')[-1]}`\n")
print("Captured with `ncu --set full --clock-control none --import-source on` on one B200; each kernel is replayed in isolation,")
print("so durations are larger than inside the CUDA graph (see profiles/README.md for in-graph times).\n")
print("| kernel | " + " | ".join(n for _, n in METRICS) + " |")
print("|---|" + "---|" * len(METRICS))
seen = set()
for d in data:
    name = d[ix["Kernel Name"]].replace("b200cnn::<unnamed>::", "").replace("void ", "")
    name = name.split("(CUtensorMap")[0].split("(const")[0][:70]
    if name in seen:
        continue
    seen.add(name)
    cells = []
    for m, _ in METRICS:
        if m in ix:
            v = d[ix[m]]
            try:
                v = f"{float(v.replace(',', '')):.4g}"
            except ValueError:
                pass
            cells.append(f"{v} {units[ix[m]]}".strip())
        else:
            cells.append("-")
    print(f"| `{name}` | " + " | ".join(cells) + " |")

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True).stdout
kernels, cur = [], None
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        kernels.append(cur)
    elif cur is not None and r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] and len(r) >= 5:
        cur["rows"].append(r)
print("\n## Most-sampled instructions (warp-state sampling, all samples)\n")
seen = set()
for k in kernels:
    if k["name"] in seen or not k["hdr"]:
        continue
    seen.add(k["name"])
    kx = {h: i for i, h in enumerate(k["hdr"])}
    s = kx["Warp Stall Sampling (All Samples)"]
    tot = sum(float(r[s] or 0) for r in k["rows"])
    short = k["name"].replace("b200cnn::<unnamed>::", "").replace("void ", "").split("(CUtensorMap")[0][:80]
    print(f"### `{short}` — {tot:.0f} samples\n\n```")
    for r in sorted(k["rows"], key=lambda r: -float(r[s] or 0))[:10]:
        print(f"{float(r[s]):7.0f} {100 * float(r[s]) / max(tot, 1):5.1f}%  {r[1].strip()[:100]}")
    print("```\n")
