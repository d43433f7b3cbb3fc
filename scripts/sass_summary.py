#!/usr/bin/env python
"""Opcode counts per kernel of the shipped libaugcl_b200.so (cuobjdump -sass): the SASS evidence for the bulk-copy /
mbarrier / packed-FMA claims in DESIGN.md.    python scripts/sass_summary.py > profiles/sass_summary.txt"""
import collections, os, re, subprocess, sys
SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "augcl_b200", "libaugcl_b200.so")
WANT = ["UBLKCP", "SYNCS", "ELECT", "ACQBULK", "UTMALDG", "STG.E.128", "LDG.E.128", "LDS.128", "STS.128", "STS.64", "FFMA2", "FFMA", "FADD", "MUFU", "PRMT",
        "HMMA", "UTCHMMA", "BAR.SYNC", "ATOMG", "LDL", "STL"]
out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
cur, per, n = None, collections.OrderedDict(), {}
for l in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", l)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        per[cur] = collections.Counter(); n[cur] = 0
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", l)
    if m and cur:
        n[cur] += 1
        op = m.group(1)
        for w in WANT:
            if op == w or op.startswith(w + "."):
                per[cur][w] += 1
print("# cuobjdump -sass of augcl_b200/libaugcl_b200.so (sm_100a): instruction count and selected opcodes per kernel")
print("# no UTMALDG / tensor-core opcode anywhere: 1-D bulk copies (UBLKCP) of contiguous planes, nothing here is a contraction")
tot = collections.Counter()
for k, c in per.items():
    tot.update(c)
    print(f"{k[:110]:110s} n={n[k]:5d} " + " ".join(f"{w}={c[w]}" for w in WANT if c[w]))
print("TOTAL " + " ".join(f"{w}={tot[w]}" for w in WANT))
