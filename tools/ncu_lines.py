"""Aggregate an `ncu --page source --csv` (SASS view) dump per CUDA source line.

usage: ncu_lines.py <sass.csv from ncu> <nvdisasm -g -c output> <mangled-function-substring> [top]

The ncu CSV lists one row per SASS instruction (address order); nvdisasm -g prints the same instructions with
`//## File "...", line N` markers (and `inlined at` chains).  Rows are matched by instruction offset.
"""
import csv
import re
import sys
from collections import defaultdict


def load_lines(sass_path, func):
    lines = {}
    cur = None
    inside = False
    with open(sass_path) as f:
        for ln in f:
            if ln.startswith(".text."):
                inside = func in ln
                continue
            if not inside:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
            if m:
                cur = (m.group(1).split("/")[-1], int(m.group(2)), m.group(3).strip())
                continue
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*)", ln)
            if m:
                lines[int(m.group(1), 16)] = cur
    return lines


def main():
    csv_path, sass_path, func = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    lines = load_lines(sass_path, func)
    rows = list(csv.reader(open(csv_path)))
    hdr = rows[1]
    ia, isamp, iex = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    base = None
    per = defaultdict(lambda: [0, 0])
    tot = [0, 0]
    for r in rows[2:]:
        if len(r) <= iex or not r[ia].startswith("0x"):
            continue
        a = int(r[ia], 16)
        if base is None:
            base = a
        key = lines.get(a - base)
        s, e = int(r[isamp] or 0), int(r[iex] or 0)
        per[key][0] += s
        per[key][1] += e
        tot[0] += s
        tot[1] += e
    print("total samples %d, warp instructions %d" % tuple(tot))
    for key, (s, e) in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%6.2f%% samples %6.2f%% inst  %s" % (100.0 * s / max(tot[0], 1), 100.0 * e / max(tot[1], 1), key))


if __name__ == "__main__":
    main()
