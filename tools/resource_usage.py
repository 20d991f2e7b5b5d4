"""Registers / stack / static shared / local memory of every kernel of the built library, as CSV:
    python tools/resource_usage.py [lib.so] > profiles/rN/resource_usage_vX.csv
(cuobjdump --dump-resource-usage; runs without a GPU).  Non-zero local bytes = spills."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "conv-tasnet_b200", "libctn_b200.so")
txt = subprocess.run(["cuobjdump", "--dump-resource-usage", lib], capture_output=True, text=True).stdout
rows, name = [], None
for line in txt.splitlines():
    m = re.match(r"\s*Function (\S+):", line)
    if m:
        name = m.group(1)
        continue
    m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+) CONSTANT\[0\]:(\d+)", line)
    if m and name:
        rows.append((name,) + tuple(int(x) for x in m.groups()))
        name = None
dem = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True, text=True).stdout.splitlines()
print("# cuobjdump --dump-resource-usage of %s (sm_100a); shared = static only" % os.path.basename(lib))
print("kernel,registers,stack_bytes,static_shared_bytes,local_bytes,param_bytes")
for r, d in sorted(zip(rows, dem), key=lambda x: x[1]):
    d = re.sub(r"\(.*", "", d).replace("void ", "").replace(",", " ")
    print("%s,%d,%d,%d,%d,%d" % (d, r[1], r[2], r[3], r[4], r[5]))
