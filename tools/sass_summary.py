"""SASS instruction summary per kernel of libctb.so (cuobjdump -sass): the mnemonics that prove what the
kernels use (DFMA FP64 pipe, UBLKCP = TMA bulk copy, SYNCS = mbarrier, ATOMS.CAS = emulated shared FP64
atomics, DMMA = FP64 tensor core). Writes profiles/r02_sass_summary.txt."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "cantera_b200", "libctb.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.split("\n")
kern, cur = collections.OrderedDict(), None
it = iter(names)
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = next(it)
        cur = cur.replace("(anonymous namespace)::", "").replace("ctb::", "")
        cur = re.sub(r"\(.*\)$", "", cur)[:90]
        while cur in kern:
            cur += "'"
        kern[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        c = kern[cur]
        c["total"] += 1
        for key in ("DFMA", "DMUL", "DADD", "DMMA", "UBLKCP", "SYNCS", "LDS", "STS", "LDG", "STG", "BAR", "SHFL", "REDUX", "MUFU"):
            if op.split(".")[0] == key:
                c[key] += 1
        if op.startswith("ATOMS.CAS"):
            c["ATOMS.CAS"] += 1
        if op.startswith("ATOMS") and not op.startswith("ATOMS.CAS"):
            c["ATOMS"] += 1
cols = ["total", "DFMA", "DMUL", "DADD", "DMMA", "UBLKCP", "SYNCS", "BAR", "LDS", "STS", "LDG", "STG", "SHFL", "REDUX", "ATOMS.CAS", "MUFU"]
with open(os.path.join(ROOT, "profiles", "r02_sass_summary.txt"), "w") as f:
    f.write("cuobjdump -sass cantera_b200/libctb.so (sm_100a), static instruction counts per kernel\n")
    f.write("%-92s" % "kernel" + "".join("%10s" % c for c in cols) + "\n")
    for k, c in kern.items():
        f.write("%-92s" % k + "".join("%10d" % c[x] for x in cols) + "\n")
    tot = collections.Counter()
    for c in kern.values():
        tot.update(c)
    f.write("%-92s" % "ALL" + "".join("%10d" % tot[x] for x in cols) + "\n")
print(open(os.path.join(ROOT, "profiles", "r02_sass_summary.txt")).read()[:3000])
