#!/usr/bin/env python
"""Warp-stall samples of one captured kernel per SOURCE LINE: aligns the SASS listing of an .ncu-rep with
`nvdisasm -g` of the same kernel in the shipped .so (the build must be the one that was profiled).
    python scripts/ncu_lines.py gpurun_out/x.ncu-rep [N] [--so augcl_b200/libaugcl_b200.so]"""
import collections, csv, io, os, re, subprocess, sys, tempfile


def sass_lines(so, kernel_mangled_part):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
    out = []
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin"):
            continue
        dis = subprocess.run(["nvdisasm", "-gi", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur, line, on, fresh = None, None, False, True
        for l in dis.splitlines():
            if l.startswith(".text."):
                on = re.search(kernel_mangled_part, l) is not None
                cur = []
                if on:
                    out.append(cur)
                continue
            if not on:
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
            if m:
                # a chain of "inlined at" records precedes an instruction: keep the OUTERMOST call site (the last record)
                line = (os.path.basename(m.group(3) or m.group(1)), int(m.group(4) or m.group(2)))
                continue
            m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", l)
            if m:
                cur.append((m.group(1).strip(), line))
    return out[0] if out else []


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    rep = args[0]
    n = int(args[1]) if len(args) > 1 else 40
    so = "augcl_b200/libaugcl_b200.so"
    if "--so" in sys.argv:
        so = sys.argv[sys.argv.index("--so") + 1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    hdr, data = rows[h], rows[h + 1:]
    name = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    kr = list(csv.reader(io.StringIO(name)))
    kname = kr[2][kr[0].index("Kernel Name")]
    m = re.match(r"(?:void )?(?:\w+::)*(\w+)<([^>]*)>", kname)
    if m:
        base, targs = m.group(1), [a.strip() for a in m.group(2).split(",")]
        pat = "%d%sI" % (len(base), base) + "".join("L[ib]%sE" % {"true": "1", "false": "0"}.get(a, a) for a in targs) + "E"
    else:   # not a template
        base = re.match(r"(?:void )?(?:\w+::)*(\w+)", kname).group(1)
        pat = "%d%sE" % (len(base), base)
    # bools are printed as 0/1 by ncu: try both encodings
    cands = [pat]
    sl = []
    for so_pat in cands:
        sl = sass_lines(so, so_pat)
        if sl:
            break
    if not sl:   # fall back: match by instruction count
        print("kernel", kname, "pattern", pat, "not found in", so)
        sys.exit(1)
    iS, iN, iE = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    cols = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
    ix = {c: hdr.index(c) for c in cols}
    if len(sl) != len(data):
        print("WARNING: instruction count differs: ncu", len(data), "nvdisasm", len(sl), "(different build?)")
    per = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    tot = 0
    for i, r in enumerate(data[:len(sl)]):
        op_ncu = r[iS].split()[0] if not r[iS].startswith("@") else r[iS].split()[1]
        key = sl[i][1]
        smp = int(r[iN] or 0)
        tot += smp
        e = per[key]
        e[0] += smp
        e[1] += int(r[iE] or 0)
        for c in cols:
            e[2][c[6:]] += int(r[ix[c]] or 0)
    print(kname, "samples", tot)
    src_cache = {}
    def text(key):
        if key is None:
            return ""
        f, ln = key
        for d in ("augcl_b200/csrc", "."):
            pth = os.path.join(d, f)
            if os.path.isfile(pth):
                if pth not in src_cache:
                    src_cache[pth] = open(pth).read().splitlines()
                L = src_cache[pth]
                return L[ln - 1].strip()[:80] if ln - 1 < len(L) else ""
        return ""
    for key, (smp, ex, st) in sorted(per.items(), key=lambda kv: -kv[1][0])[:n]:
        top = ", ".join(f"{k}:{v}" for k, v in st.most_common(2))
        print(f"{smp:6d} {100.0 * smp / max(tot, 1):5.1f}%  exec {ex:>10d}  {str(key):28s} {top:34s} | {text(key)}")


if __name__ == "__main__":
    main()
