"""Markdown table of an all-reduce sweep JSON (bench/allreduce_sweep.py --out)."""
import collections, json, sys
rows = json.load(open(sys.argv[1]))
tab = collections.OrderedDict()
for r in rows:
    tab.setdefault(r["bytes"], {})[r["path"]] = r
paths = ["fused-sgd/auto", "fused-sgd/one_shot", "fused-sgd/two_shot", "fused-sgd/nvls", "fused-ar/nvls", "nccl", "nccl+sgd"]
w = rows[0]["world"]
print(f"world = {w}, multicast = {rows[0]['multicast']}, region = {rows[0]['region']}; cells: time in us / bus bandwidth GB/s (algbw x 2(W-1)/W)\n")
print("| bytes | " + " | ".join(paths) + " | best fused-sgd frac of 770 GB/s |")
print("|---|" + "---|" * (len(paths) + 1))
for b, d in tab.items():
    best = max((d[p]["frac_of_link"] for p in paths[:4] if p in d), default=0)
    print(f"| {b} | " + " | ".join((f"{d[p]['us']:.1f} / {d[p]['busbw_gbs']:.1f}" if p in d else "-") for p in paths) + f" | {best:.2f} |")
