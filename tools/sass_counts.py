#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA in libb2c.so
(cuobjdump -sass): UTCHMMA (tcgen05.mma), LDTM (tcgen05.ld), UTMALDG / UTMASTG (TMA load / store),
UTCBAR (tcgen05.commit), SYNCS (mbarrier).  Writes profiles/r02_sass_counts.txt.

  python tools/sass_counts.py [lib] [out]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "3d-gan_b200", "b200conv3d", "_lib", "libb2c.so")
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "r02_sass_counts.txt")
MN = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "UTMALDG", "UTMASTG", "UTCBAR", "UTCATOMSWS", "SYNCS", "HMMA", "FFMA"]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
fn, rows = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        rows.setdefault(fn, collections.Counter())
        continue
    if fn is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op = m.group(1)
    for k in MN:
        if op == k or op.startswith(k + "."):
            rows[fn][k] += 1
    if op.startswith("UTCHMMA") and ".2CTA" in op:
        rows[fn]["UTCHMMA.2CTA"] += 1
with open(out, "w") as f:
    f.write("# cuobjdump -sass %s : mnemonic counts per kernel (sm_100a)\n" % os.path.relpath(lib, ROOT))
    f.write("%-64s " % "kernel" + " ".join("%12s" % k for k in MN) + "\n")
    tot = collections.Counter()
    for k, c in rows.items():
        tot.update(c)
        if any(c[m] for m in MN[:7]):
            f.write("%-64s " % k[:64] + " ".join("%12d" % c[m] for m in MN) + "\n")
    f.write("%-64s " % ("TOTAL over %d kernels" % len(rows)) + " ".join("%12d" % tot[m] for m in MN) + "\n")
print(open(out).read())
