#!/usr/bin/env python
"""Per-source-line shares of a kernel in an `ncu --set full --import-source on` report (the CSV source page has SASS only):
the SASS of the same library build is taken from `nvdisasm -g` and matched instruction by instruction.

    python tools/ncu_lines.py REPORT.ncu-rep LIBRARY.so KERNEL_SUBSTRING [SOURCE_FILE] [min_percent]
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile


def main():
    rep, lib, kern = sys.argv[1:4]
    srcfile = sys.argv[4] if len(sys.argv) > 4 else None
    floor = float(sys.argv[5]) if len(sys.argv) > 5 else 0.8
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, check=True, capture_output=True)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    funcs, ops, cur, line = {}, {}, None, None
    for l in dis.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
        if m:
            cur = m.group(1); funcs[cur] = []; ops[cur] = []; line = None
            continue
        m = re.search(r'//## File ".*?([\w\.]+)", line (\d+)', l)
        if m:
            line = (m.group(1), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(.*?)\s*;", l)
        if cur and m:
            funcs[cur].append(line); ops[cur].append(m.group(1))
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    secs, c, hdr = [], None, None
    for r in rows:
        if r and r[0] == "Kernel Name":
            c = [r[1], []]; secs.append(c)
        elif r and r[0] == "Address":
            hdr = r
        elif r and r[0].startswith("0x"):
            c[1].append(r)
    ci, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    src = open(srcfile).read().split("\n") if srcfile else None

    def opc(t):
        return re.sub(r"^@!?U?P\d+\s+", "", t.strip()).split()[0]
    done = set()
    for name, ins in secs:
        if kern not in name or name in done:
            continue
        done.add(name)
        tag = re.sub(r"[^A-Za-z0-9]", "", name.split("(")[0].split("::")[-1])
        cands = [f for f in funcs if len(funcs[f]) == len(ins) and all(opc(a) == opc(b[1]) for a, b in zip(ops[f], ins))]
        if not cands:
            print(name, ": no function of", lib, "has this SASS (another build?)")
            continue
        fn = cands[0]
        agg, smp = collections.Counter(), collections.Counter()
        for ln, r in zip(funcs[fn], ins):
            agg[ln] += int(r[ci]); smp[ln] += int(r[si])
        tot, ts = sum(agg.values()), sum(smp.values())
        print("%s: %d warp instructions, %d samples" % (name, tot, ts))
        for ln in sorted(agg, key=lambda x: (x is None, x)):
            if smp[ln] / ts * 100 >= floor or agg[ln] / tot * 100 >= floor:
                text = src[ln[1] - 1].strip()[:110] if ln and src and ln[0] == os.path.basename(srcfile) else ""
                print("  %5.1f%% samples %5.1f%% instr  %s:%s  %s" % (100 * smp[ln] / ts, 100 * agg[ln] / tot, ln[0] if ln else "?", ln[1] if ln else "", text))


if __name__ == "__main__":
    main()
