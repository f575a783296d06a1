"""Attribute the stall samples of an `ncu --page source --csv` SASS export to CUDA source lines, using the line
table nvdisasm prints for the same cubin.  usage: ncu_lines.py <source.csv> <all.sass> <mangled kernel name> [min %]"""
import csv
import re
import sys
from collections import defaultdict


def main():
    src_csv, sass, kernel = sys.argv[1:4]
    min_pct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.7
    lines = open(sass).read().split("\n")
    start = [i for i, l in enumerate(lines) if l.startswith(".text.%s:" % kernel)][0]
    off2line, cur = {}, None
    for l in lines[start + 1:]:
        if l.startswith("//---------------------"):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            off2line[int(m.group(1), 16)] = cur
    rows = list(csv.reader(open(src_csv)))
    hdr = rows[1]
    si, ie = hdr.index("# Samples"), hdr.index("Instructions Executed")
    data = rows[2:]
    base = int(data[0][0], 16)
    agg, tot = defaultdict(lambda: [0, 0]), 0
    for r in data:
        c = off2line.get(int(r[0], 16) - base)
        agg[c][0] += int(r[si])
        agg[c][1] += int(r[ie])
        tot += int(r[si])
    cache = {}
    print("total samples", tot)
    for c, (s, e) in sorted(agg.items(), key=lambda kv: (kv[0] or ("", 0))):
        if s >= tot * min_pct / 100:
            f, ln = c if c else ("?", 0)
            if f not in cache:
                try:
                    cache[f] = open(f).read().split("\n")
                except OSError:
                    cache[f] = []
            text = cache[f][ln - 1].strip()[:100] if 0 < ln <= len(cache[f]) else ""
            print("%-18s %4d %5.1f%% instr=%10d  %s" % (f.split("/")[-1], ln, 100 * s / tot, e, text))


if __name__ == "__main__":
    main()
