#!/usr/bin/env python
"""profiles/sass_summary.txt: per kernel of libmiraculix_b200.so, the SASS mnemonics that prove the
Blackwell-native paths (B200_PROFILING.md): UTC*MMA = tcgen05.mma, UTMALDG = TMA loads, LDTM / STTM =
tcgen05.ld / st, UTCBAR = tcgen05.commit, SYNCS = mbarrier, FENCE.VIEW.ASYNC = fence.proxy.async, and
LD.E / ST.E on what are peer pointers in the exchange kernels.  Run here (no GPU): cuobjdump -sass."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "miraculix_b200", "libmiraculix_b200.so")
PAT = ["UTCIMMA", "UTCIMMA.2CTA", "UTMALDG", "UTMALDG.2D.2CTA", "LDTM", "STTM", "UTCBAR", "UTCBAR.2CTA",
       "SYNCS", "FENCE.VIEW.ASYNC", "ELECT", "LDS", "STS", "LDG", "LD.E", "STG", "ST.E", "RED", "ATOM", "DFMA", "IDP"]
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
counts, order, cur = collections.defaultdict(collections.Counter), [], None
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur)
        order.append(cur)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
    if m and cur:
        op = m.group(1)
        for p in PAT:
            if op == p or op.startswith(p + "."):
                counts[cur][p] += 1
print("# cuobjdump -sass miraculix_b200/libmiraculix_b200.so | per-kernel counts of the mnemonics that matter")
print("# (scripts/sass_summary.py; a prefix also counts its suffixed forms, e.g. UTCIMMA includes UTCIMMA.2CTA)")
for k in order:
    c = counts[k]
    if not c:
        continue
    print(f"{k}\n    " + "  ".join(f"{p}={c[p]}" for p in PAT if c[p]))
