"""Per-kernel counts of the Blackwell-native SASS mnemonics in the built library (cuobjdump -sass):
UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UBLKCP = TMA, UTCBAR = tcgen05.commit,
SYNCS = mbarrier, plus a short excerpt around the first MMA of each tensor kernel.
    python tools/sass_evidence.py > profiles/r2/sass_tensor_kernels.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "conv-tasnet_b200", "libctn_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
PAT = re.compile(r"\b(UTC[A-Z]*MMA(?:\.[A-Z0-9_]+)*|LDTM(?:\.[A-Z0-9x_]+)*|STTM[.\w]*|UTMALDG(?:\.[A-Z0-9_]+)*|UTMASTG[.\w]*|"
                 r"UBLKCP[.\w]*|UTCBAR(?:\.[A-Z0-9_]+)*|UTCATOMSWS[.\w]*|HMMA[.\w]*|FFMA2|FMUL2|FADD2)\b")
fn, counts, first = None, collections.OrderedDict(), {}
lines = sass.splitlines()
for i, ln in enumerate(lines):
    m = re.search(r"Function : (\S+)", ln)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[fn] = collections.Counter()
        continue
    if fn is None:
        continue
    for t in PAT.findall(ln):
        counts[fn][t] += 1
        if t.startswith("UTC") and "MMA" in t and fn not in first:
            first[fn] = i
print("# cuobjdump -sass conv-tasnet_b200/libctn_b200.so (sm_100a): Blackwell-native instructions per kernel")
for fn, c in counts.items():
    if any(k.startswith(("UTC", "LDTM", "UTMA", "UBLK")) for k in c):
        print("\n%s" % fn)
        for k, v in sorted(c.items()):
            print("    %-40s %d" % (k, v))
print("\n# excerpts: the first tcgen05.mma of each tensor kernel and its neighbourhood")
for fn, i in first.items():
    print("\n== %s" % fn)
    for ln in lines[max(0, i - 6):i + 8]:
        print("   " + re.sub(r"/\*[0-9a-f]{4}\*/", "", ln).strip()[:150])
