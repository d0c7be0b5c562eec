"""Aggregate the warp-stall samples of an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line.
usage: python tools/ncu_source_lines.py dump.csv file_substring [top_n]"""
import collections
import csv
import sys

path, want = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
rows = list(csv.reader(open(path)))
cur, hdr, last = None, None, None
agg = collections.defaultdict(collections.Counter)
tot = 0
lines = {}
for r in rows:
    if r and r[0] == "File Path":
        cur, hdr = r[1], None
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if not hdr or not r or want not in (cur or ""):
        continue
    if r[0]:                               # a CUDA source line (its text may hold unquoted commas: only take the number)
        last = r[0]
        lines.setdefault(last, ",".join(r[1:4]))
        continue
    if len(r) != len(hdr) or last is None:
        continue
    d = dict(zip(hdr, r))
    def num(k):
        try:
            return int(d.get(k) or 0)
        except ValueError:
            return 0
    n = num("# Samples")
    agg[last]["n"] += n
    agg[last]["inst"] += num("Instructions Executed")
    tot += n
    for k in hdr:
        if k.startswith("stall_") and "Not Issued" not in k:
            agg[last][k] += num(k)
print("total samples", tot)
for ln, c in sorted(agg.items(), key=lambda kv: -kv[1]["n"])[:top]:
    st = sorted(((k[6:], v) for k, v in c.items() if k.startswith("stall_")), key=lambda kv: -kv[1])[:3]
    print(f"{ln:>5} {c['n']:7d} {100 * c['n'] / max(tot, 1):5.1f}%  inst {c['inst']:10d}  {st}  | {lines.get(ln, '')[:60].strip()}")
