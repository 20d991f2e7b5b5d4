"""SASS listing with stall samples: python tools/ncu_sass.py rep.ncu-rep [grep-pattern context]"""
import csv, io, subprocess, sys, re
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
h = rows[hi[0]]
ci = {n: i for i, n in enumerate(h)}
body = [r for r in rows[hi[0] + 1:(hi[1] - 1 if len(hi) > 1 else None)] if len(r) > 10]
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 12
keep = set()
for i, r in enumerate(body):
    if pat is None or pat.search(r[ci["Source"]]):
        keep.update(range(max(0, i - ctx), min(len(body), i + ctx + 1)))
last = -2
for i in sorted(keep):
    if i != last + 1:
        print("   ...")
    last = i
    r = body[i]
    st = sorted(((n[6:], int(r[ci[n]] or 0)) for n in stalls), key=lambda x: -x[1])[:1]
    print("%4d %5s %7s  %-80s %s" % (i, r[ci["# Samples"]], r[ci["Instructions Executed"]], r[ci["Source"]].strip()[:80], st[0] if st[0][1] else ""))
