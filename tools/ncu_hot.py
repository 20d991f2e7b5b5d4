"""Top stalled SASS instructions of an ncu report: python tools/ncu_hot.py rep.ncu-rep [N]"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
import os
KI = int(os.environ.get("NCU_KERNEL", "0"))      # which kernel of the report (NCU_KERNEL=1: the second launch captured)
h = rows[hi[KI]]
ci = {n: i for i, n in enumerate(h)}
body = [r for r in rows[hi[KI] + 1:(hi[KI + 1] - 1 if len(hi) > KI + 1 else None)] if len(r) > 10]
tot = sum(int(r[ci["# Samples"]] or 0) for r in body)
print("total samples", tot, "instructions", len(body))
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
agg = {n: sum(int(r[ci[n]] or 0) for r in body) for n in stalls}
print("stall totals:", sorted(agg.items(), key=lambda x: -x[1])[:8])
N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
for r in sorted(body, key=lambda r: -int(r[ci["# Samples"]] or 0))[:N]:
    st = sorted(((n, int(r[ci[n]] or 0)) for n in stalls), key=lambda x: -x[1])[:2]
    print(r[ci["# Samples"]].rjust(6), r[ci["Instructions Executed"]].rjust(8), r[ci["Source"]].strip()[:72].ljust(72), st)
