"""Per-kernel totals and shares of an ncu launch list (the `--metrics gpu__time_duration.sum --csv` pass):
python tools/launch_shares.py profiles/r1_launches_ncu_522.csv"""
import collections
import csv
import re
import sys

rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
rd = csv.DictReader(rows)
tot = collections.defaultdict(lambda: [0, 0.0])
for r in rd:
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    name = re.sub(r"^void ", "", name)
    name = name.replace("gsr::<unnamed>::", "")[:70]
    ns = float(r["Metric Value"].replace(",", ""))
    tot[name][0] += 1
    tot[name][1] += ns
# the bench also times the reference CUDA (oracle/_ref) and uses a few torch fills: not product kernels
OTHER = ("renderCUDA", "preprocessCUDA", "cub::", "duplicateWithKeys", "identifyTileRanges", "computeCov2DCUDA", "at::",
         "checkFrustum")
mine = {k: v for k, v in tot.items() if not k.startswith(OTHER)}
allns = sum(v[1] for v in mine.values())
print("product kernels (share of the product's own GPU time)")
print(f"{'kernel':70s} {'launches':>8s} {'total us':>10s} {'avg us':>9s} {'share':>6s}")
for k, (n, ns) in sorted(mine.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:70s} {n:8d} {ns / 1e3:10.1f} {ns / 1e3 / n:9.2f} {ns / allns:6.3f}")
print("\nother launches in the same command (reference CUDA baseline leg, torch fills)")
for k, (n, ns) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    if k not in mine:
        print(f"{k:70s} {n:8d} {ns / 1e3:10.1f} {ns / 1e3 / n:9.2f}")
