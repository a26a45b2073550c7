#!/usr/bin/env python
"""Per-source-line hot spots of an .ncu-rep captured with --import-source on (no GPU needed).

    python tools/ncu_lines.py prof.ncu-rep [kernel-substring] [top]

ncu's CSV source page is per SASS instruction; this joins it, by instruction order, with `nvdisasm -g` line info of the
cubins inside docodo_b200/libdocodo_b200.so and sums warp-instructions and stall samples per source line.
"""
import collections
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line_tables():
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "docodo_b200", "libdocodo_b200.so")], cwd=tmp, capture_output=True)
    tables = {}
    for cub in glob.glob(os.path.join(tmp, "*.cubin")):
        out = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
        fn, cur, lst = None, None, None
        for ln in out.splitlines():
            m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
            if m:
                fn = m.group(1)
                lst = tables.setdefault(fn, [])
                cur = None
                continue
            m = re.search(r'//## File "(.*)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            if fn and re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
                lst.append(cur)
    return tables


def demangle(names):
    out = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    rep = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    tables = line_tables()
    dem = demangle(list(tables))
    by_dem = {re.sub(r"\s+", "", v): k for k, v in dem.items()}
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    i = 0
    seen = set()
    while i < len(rows):
        if rows[i] and rows[i][0] == "Kernel Name":
            kname = rows[i][1]
            hdr = rows[i + 1]
            j = i + 2
            body = []
            while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
                body.append(rows[j])
                j += 1
            i = j
            if want not in kname or kname in seen:
                continue
            seen.add(kname)
            key = by_dem.get(re.sub(r"\s+", "", kname.replace("(int)", "").replace("(bool)", "")))
            if key is None:   # try loose match on the function name
                base = re.sub(r"^void ", "", kname).split("(")[0].split("<")[0].split("::")[-1]
                cands = [k for k, v in dem.items() if base in v]
                key = cands[0] if len(cands) == 1 else None
                if key is None and cands:
                    targ = re.sub(r"[^0-9]", "", kname.split("(dcd")[0].split("(const")[0])
                    for c_ in cands:
                        if re.sub(r"[^0-9]", "", dem[c_].split("(dcd")[0].split("(const")[0]) == targ:
                            key = c_
            lines = tables.get(key, [])
            col = {h: n for n, h in enumerate(hdr)}
            ci, cs = col["Instructions Executed"], col["Warp Stall Sampling (All Samples)"]
            stall_cols = [(h, n) for h, n in col.items() if h.startswith("stall_") and "Not Issued" not in h]
            agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
            tot_i = tot_s = 0
            for n, r in enumerate(body):
                if len(r) <= cs:
                    continue
                loc = lines[n] if n < len(lines) else None
                a = agg[loc]
                ins, smp = int(r[ci] or 0), int(r[cs] or 0)
                a[0] += ins
                a[1] += smp
                tot_i += ins
                tot_s += smp
                for h, c in stall_cols:
                    v = int(r[c] or 0)
                    if v:
                        a[2][h[6:]] += v
            print("=== %s  (warp-inst %d, samples %d, sass rows %d, line-table %d)" % (kname[:100], tot_i, tot_s, len(body), len(lines)))
            for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
                st = " ".join("%s:%d" % (k, v) for k, v in a[2].most_common(3))
                print("  %-28s inst %5.1f%%  samples %5.1f%%  %s" % ("%s:%s" % loc if loc else "?", 100.0 * a[0] / max(tot_i, 1),
                                                                  100.0 * a[1] / max(tot_s, 1), st))
        else:
            i += 1


if __name__ == "__main__":
    main()
