"""Per-SASS-instruction executed counts of one kernel from an .ncu-rep: python tools/ncu_sass.py rep regex [min_pct]"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.3
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = [i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r][0]
H = rows[hi]
ia, ie, isamp, ite = H.index("Source"), H.index("Instructions Executed"), H.index("# Samples"), H.index("Avg. Threads Executed")
body = []
for r in rows[hi + 1:]:
    if len(r) > ie and r[ie].isdigit():
        body.append(r)
    elif "Source" in r and "Instructions Executed" in r:
        break
tot = sum(int(r[ie]) for r in body)
print("total warp instructions", tot, "SASS lines", len(body))
for n, r in enumerate(body):
    e = int(r[ie])
    if e >= tot * minpct / 100:
        print(f"{n:4d} {e / 1e6:8.2f}M {100 * e / tot:5.1f}% thr={r[ite]:>5s} samp={r[isamp]:>6s}  {r[ia].strip()[:100]}")
