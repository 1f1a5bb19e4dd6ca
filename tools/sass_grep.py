"""Counts of the Blackwell-native SASS mnemonics per kernel of libcpoly.so (no GPU needed):
UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UBLKCP = cp.async.bulk, UTMALDG / UTMASTG = TMA tensor copies,
HMMA = legacy mma.sync, LDGSTS = cp.async, RED / ATOM(G|S) = atomics.   python tools/sass_grep.py > profiles/rNN_sass_grep.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "centerpoly_b200", "libcpoly.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = collections.OrderedDict([("UTC*MMA", r"\bUTC\w*MMA\b"), ("LDTM", r"\bLDTM\b"), ("STTM", r"\bSTTM\b"), ("UBLKCP", r"\bUBLKCP\b"),
                                ("UTMALDG", r"\bUTMALDG\b"), ("UTMASTG", r"\bUTMASTG\b"), ("HMMA", r"\bHMMA\b"), ("LDGSTS", r"\bLDGSTS\b"),
                                ("REDG.F32", r"\bREDG\.E\.ADD\.F32"), ("ATOMG", r"\bATOMG\b"), ("ATOMS", r"\bATOMS\b"), ("MATCH", r"\bMATCH\b")])
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        k = re.search(r"(?:<unnamed>::|cpoly::)+(\w+)(<[^(]*>)?", name)
        cur = (k.group(1) + (k.group(2) or "")) if k else name[:80]
        counts[cur] = collections.Counter()
        continue
    if cur:
        for key, pat in pats.items():
            if re.search(pat, line):
                counts[cur][key] += 1
print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}: instruction counts per kernel (sm_100a)")
print("%-62s" % "kernel" + "".join("%9s" % k for k in pats))
tot = collections.Counter()
for name, c in counts.items():
    if sum(c.values()) == 0:
        continue
    print("%-62s" % name[:62] + "".join("%9d" % c[k] for k in pats))
    tot.update(c)
print("%-62s" % "TOTAL" + "".join("%9d" % tot[k] for k in pats))
