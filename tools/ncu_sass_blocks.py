"""Executed warp-instructions of one kernel grouped into straight-line SASS runs with equal counts:
python tools/ncu_sass_blocks.py rep regex"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = [i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r][0]
H = rows[hi]
ia, ie, ite = H.index("Source"), H.index("Instructions Executed"), H.index("Avg. Threads Executed")
body = []
for r in rows[hi + 1:]:
    if len(r) > ie and r[ie].isdigit():
        body.append(r)
    elif "Source" in r and "Instructions Executed" in r:
        break
tot = sum(int(r[ie]) for r in body)
print("total warp instructions", tot, "SASS lines", len(body))
i = 0
while i < len(body):
    j = i
    e = int(body[i][ie])
    while j + 1 < len(body) and abs(int(body[j + 1][ie]) - e) <= 0.02 * max(e, 1):
        j += 1
    s = sum(int(r[ie]) for r in body[i:j + 1])
    if s >= tot * 0.004:
        ops = " ".join(r[ia].split()[0 if not r[ia].strip().startswith("@") else 1] for r in body[i:j + 1])[:110]
        print(f"{i:4d}-{j:4d} n={j - i + 1:3d} each={e / 1e6:7.2f}M sum={100 * s / tot:5.1f}% thr={body[i][ite]:>4s} {ops}")
    i = j + 1
