#!/usr/bin/env python
"""Hottest SASS instructions of an .ncu-rep by warp-stall samples, with their dominant stall reasons.
    python scripts/ncu_hot.py gpurun_out/x.ncu-rep [N]"""
import csv, io, subprocess, sys

def main(rep, n):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    hdr, data = rows[h], rows[h + 1:]
    iS, iN, iE = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    cols = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
    ix = {c: hdr.index(c) for c in cols}
    tot = sum(int(r[iN] or 0) for r in data)
    print("total samples", tot, "instructions", len(data))
    print({c[6:]: sum(int(r[ix[c]] or 0) for r in data) for c in cols if sum(int(r[ix[c]] or 0) for r in data)})
    top = sorted(range(len(data)), key=lambda i: -int(data[i][iN] or 0))[:n]
    for i in sorted(top):
        r = data[i]
        st = sorted(((c[6:], int(r[ix[c]] or 0)) for c in cols), key=lambda kv: -kv[1])[:2]
        print(f"{i:5d} {int(r[iN] or 0):6d} {r[iE]:>9s}  {r[iS][:64]:64s} {st}")

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
