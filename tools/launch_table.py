"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` log: mean device time per distinct kernel."""
import collections, csv, sys
rows = list(csv.DictReader(l for l in open(sys.argv[1]) if not l.startswith("==")))
agg = collections.OrderedDict()
for r in rows:
    k = r["Kernel Name"].replace("b200cnn::<unnamed>::", "").replace("void ", "")[:64] + " grid=" + r.get("Grid Size", "") + " blk=" + r.get("Block Size", "")
    agg.setdefault(k, []).append(float(r["Metric Value"].replace(",", "")))
tot = 0
print("| mean us | launches | kernel |\n|---|---|---|")
for k, v in agg.items():
    print(f"| {sum(v)/len(v)/1000:7.2f} | {len(v)} | `{k}` |")
    tot += sum(v) / 1000
print(f"\ntotal {tot:.1f} us over {len(rows)} launches")
